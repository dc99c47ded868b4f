// What every plugin translation unit defines for itself: the error text behind its last_error() and the SM count.
#pragma once
#include <cstdarg>
#include "oz_common.cuh"

namespace oz {
static thread_local char g_plugin_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_plugin_err, sizeof(g_plugin_err), fmt, ap);
    va_end(ap);
}
int sm_count() {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
        return 148;
    return n;
}
}  // namespace oz
