// FP64 FMA throughput probe: the denominator for the FP64-bound paths (implicit Gauss-Legendre + Newton, the
// exp-heavy grid step).  BASELINE.md asks for the FP64 vector peak to be MEASURED on the box because config C2
// is bound by it, not by HBM.  16 independent DFMA chains per thread, 8 resident blocks of 256 threads per SM.
#include "oz_common.cuh"

namespace {
template <typename T>
__global__ void __launch_bounds__(256) fma_probe_kernel(T* out, int iters, T seed) {
    T acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = seed + T(i) * T(1e-3) + T(threadIdx.x) * T(1e-6);
    const T a = T(1.0000001), b = T(1e-9);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = acc[i] * a + b;
    }
    T s = T(0);
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    if (s == T(-1)) out[0] = s;  // never true: keeps the chains alive
}
}  // namespace

extern "C" {
// Enqueue the probe on `stream`; returns the number of floating-point operations it performs (2 per FMA).
// The caller times it with events.  dtype: OZ_F64 / OZ_F32.  `out` is one scalar of device memory.
OZ_API double oz_fma_probe(int dtype, int iters, void* out, void* stream) {
    const int blocks = oz::sm_count() * 8;
    if (dtype == OZ_F64)
        fma_probe_kernel<double><<<blocks, 256, 0, (cudaStream_t)stream>>>((double*)out, iters, 1.0);
    else
        fma_probe_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>((float*)out, iters, 1.0f);
    return 2.0 * 16.0 * (double)iters * 256.0 * (double)blocks;
}
}
