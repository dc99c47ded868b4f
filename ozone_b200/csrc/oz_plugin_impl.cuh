// Included by a generated plugin translation unit AFTER its functor definition:
//     struct RhsTraced { ... };           (same contract as the functors in oz_rhs.cuh)
//     #include "oz_plugin_impl.cuh"
// Only the (stage count, explicit/implicit, dtype) combination named by the OZ_ONLY_* macros is
// instantiated, so a plugin compiles in seconds.
#pragma once
#include <cstdarg>
#include "oz_plugin.h"
#include "oz_tm_launch.cuh"
#include "oz_plugin_common.cuh"

OZ_INSTANTIATE_TM(oz::RhsTraced, plugin)

extern "C" __attribute__((visibility("default"))) void oz_plugin_describe(oz_plugin_info* out) {
    out->abi = OZ_PLUGIN_ABI;
    out->d = oz::RhsTraced::D;
    out->P = oz::RhsTraced::P;
    out->S = oz::RhsTraced::S;
    out->nprof = oz::RhsTraced::NPROF;
    out->nfield = oz::RhsTraced::NFIELD;
    out->tm = oz::tm_plugin;
    out->picard = oz::picard_plugin;
    out->colloc = oz::colloc_plugin;
    out->last_error = []() -> const char* { return oz::g_plugin_err; };
    out->kind = 0;
}
