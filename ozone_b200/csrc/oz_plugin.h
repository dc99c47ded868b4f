// Interface between liboz_b200.so and a separately compiled right-hand side (see ozone_b200/trace.py).
// A plugin is one .cu file: a generated functor + OZ_INSTANTIATE_TM(functor, plugin) + OZ_DEFINE_PLUGIN.
#pragma once
#include <stddef.h>
#include "../../include/ozone_b200.h"

#define OZ_PLUGIN_ABI 1

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
} oz_plugin_info;
