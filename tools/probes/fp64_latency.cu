// Dependent-issue latency and per-scheduler throughput of the FP64 pipe (DFMA), measured with clock64().
//   chains = independent DFMA chains per thread (ILP), warps = warps per CTA (one CTA on one SM).
// Prints cycles per DFMA per warp.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/fp64_latency tools/probes/fp64_latency.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int CHAINS>
__global__ void probe(double* out, long long* cyc, int iters, double a, double b) {
    double x[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) x[c] = threadIdx.x * 1e-3 + c;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < CHAINS; ++c) x[c] = fma(x[c], a, b);
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int CHAINS>
void run(int warps, double* out, long long* cyc) {
    const int iters = 2000;
    probe<CHAINS><<<1, 32 * warps>>>(out, cyc, iters, 0.999999, 1e-9);
    cudaDeviceSynchronize();
    probe<CHAINS><<<1, 32 * warps>>>(out, cyc, iters, 0.999999, 1e-9);
    cudaDeviceSynchronize();
    long long c;
    cudaMemcpy(&c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
    const double per = (double)c / (iters * 8.0 * CHAINS);
    printf("warps/CTA %2d (per scheduler %.2f)  chains %d : %.2f cycles per DFMA per warp, %.3f DFMA/cycle/scheduler\n", warps,
           warps / 4.0, CHAINS, per, (warps / 4.0) / per);
}

int main() {
    double* out;
    long long* cyc;
    cudaMalloc(&out, 1024 * sizeof(double));
    cudaMalloc(&cyc, 64 * sizeof(long long));
    for (int w : {1, 4, 8, 12, 16}) {
        run<1>(w, out, cyc);
        run<2>(w, out, cyc);
        run<4>(w, out, cyc);
        run<8>(w, out, cyc);
    }
    return 0;
}
