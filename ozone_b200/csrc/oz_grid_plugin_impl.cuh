// Included by a generated grid plugin AFTER its model (ozone_b200/trace_grid.py):
//     #include "oz_pde_kernels.cuh"
//     namespace oz { template <typename T_> struct GridTraced { ... }; }      (the model contract of oz_pde_kernels.cuh)
//     #include "oz_grid_plugin_impl.cuh"
// Only the (stage count, dtype) named by the OZ_ONLY_* macros is instantiated.
#pragma once
#include <cstdarg>
#include "oz_plugin.h"
#include "oz_plugin_common.cuh"

namespace oz {
static int grid_plugin_tm(const oz_problem* pb, int mode, size_t* ws_bytes, const EvalArgs* ev) {
    if (int rc = pde_check(pb, mode, ev)) return rc;
    if constexpr (OZ_WANT_F64(true))
        if (pb->dtype == OZ_F64) return pde_by_stage<GridTraced<double>>(pb, mode, ws_bytes);
    if constexpr (OZ_WANT_F64(false))
        if (pb->dtype == OZ_F32) return pde_by_stage<GridTraced<float>>(pb, mode, ws_bytes);
    set_error("no grid kernel for dtype %d in this build", pb->dtype);
    return -1;
}
static int grid_plugin_no_picard(const oz_problem*, const void*, double, int, const int*, int) {
    set_error("a grid right-hand side runs under TimeMarching / TimeMarchingCheckpoints only");
    return -1;
}
static int grid_plugin_no_colloc(const oz_problem*, int, size_t*, const void*, const void*, const void*, const void*, void*, void*) {
    set_error("a grid right-hand side runs under TimeMarching / TimeMarchingCheckpoints only");
    return -1;
}
}  // namespace oz

extern "C" __attribute__((visibility("default"))) void oz_plugin_describe(oz_plugin_info* out) {
    out->abi = OZ_PLUGIN_ABI;
    out->d = 1;
    out->P = oz::GridTraced<double>::P;
    out->S = oz::GridTraced<double>::NCOEF;
    out->nprof = 0;
    out->nfield = oz::GridTraced<double>::NFIELD;
    out->tm = oz::grid_plugin_tm;
    out->picard = oz::grid_plugin_no_picard;
    out->colloc = oz::grid_plugin_no_colloc;
    out->last_error = []() -> const char* { return oz::g_plugin_err; };
    out->kind = 1;
}
