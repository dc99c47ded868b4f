// Interface between liboz_b200.so and a separately compiled right-hand side (see ozone_b200/trace.py).
// A plugin is one generated .cu file: a functor followed by oz_plugin_impl.cuh (per-trajectory right-hand sides) or a grid
// model followed by oz_grid_plugin_impl.cuh (one trajectory whose state is a 1-D grid, see oz_pde_kernels.cuh).
#pragma once
#include <stddef.h>
#include "../../include/ozone_b200.h"

#define OZ_PLUGIN_ABI 2

namespace oz {
struct EvalArgs;
}

typedef struct oz_plugin_info {
    int abi;
    int d, P, S, nprof, nfield;
    int (*tm)(const oz_problem*, int mode, size_t* ws_bytes, const oz::EvalArgs*);
    int (*picard)(const oz_problem*, const void* guess, double tol, int max_sweeps, const int* blk_lo, int nblocks);
    int (*colloc)(const oz_problem*, int mode, size_t* ws_bytes, const void* Y, const void* y, const void* mu1,
                  const void* mu2, void* o1, void* o2);
    const char* (*last_error)(void);
    int kind;   // 0: one ODE per trajectory (oz_plugin_impl.cuh); 1: a grid right-hand side (oz_grid_plugin_impl.cuh): only `tm` is used,
                // d is a placeholder (the cell count comes with each problem as d_override), S counts coefficient FIELDS
} oz_plugin_info;
