// Neighbour halo exchange and small all-reduce over peer-mapped memory (NVLink / NVSwitch), as plain kernels.
//
// The reference has no distributed code; these entry points exist because the two collectives on the
// multi-GPU paths of this engine are tiny and latency-bound, and an NCCL call per segment (a host-blocking
// send/recv pair, ~20-40 us) costs more than the kernels it separates:
//   * the spatial shard of the diffusion-reaction grid exchanges 2 s n_inner cells per side once per
//     checkpoint segment (ozone_b200/parallel.py);
//   * the trajectory-batch shard sums the gradients of batch-shared inputs (a few dozen scalars).
// Every rank owns an "arena" allocated in symmetric memory, so each rank holds a device pointer to every
// other rank's arena.  Data moves with ordinary stores into the peer's arena followed by a release store
// of an epoch number into a flag in the same arena; the receiver polls its OWN memory (acquire loads) and
// then reads the payload through L2 (ld.global.cg: peer writes are coherent at L2, not in L1).
// Payload slots are double-buffered by epoch parity; a rank can only run two epochs ahead of a neighbour
// after that neighbour has signalled the epoch in between, which it does after consuming the older slot.
// Epochs are `*epoch_base + epoch`, epoch_base living in device memory, so a captured CUDA graph can be
// replayed: oz_peer_epoch_advance moves the base at the end of every run.
// A wait gives up after ~2 s of polling (clock64) and raises the error flag instead of hanging the GPU.
// No TMA, no multicast: loads and stores only.
#include "oz_common.cuh"

namespace oz {

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

constexpr long long kPeerTimeoutCycles = 4000000000ll;   // ~2 s at 2 GHz

// true when *flag reached `target` before the timeout
__device__ bool peer_wait(const unsigned long long* flag, unsigned long long target) {
    const long long t0 = clock64();
    while (ld_acquire_sys(flag) < target) {
        if (clock64() - t0 > kPeerTimeoutCycles) return false;
        __nanosleep(64);
    }
    return true;
}

// Copy `count` values to the left and/or right neighbour's slot, then publish the epoch.  One CTA.
template <typename T>
__global__ void __launch_bounds__(256) halo_push_kernel(const T* __restrict__ left_src, const T* __restrict__ right_src,
                                                        long long count, T* left_dst, T* right_dst,
                                                        unsigned long long* left_flag, unsigned long long* right_flag,
                                                        const unsigned long long* epoch_base, unsigned long long epoch) {
    for (long long j = threadIdx.x; j < count; j += blockDim.x) {
        if (left_dst) left_dst[j] = left_src[j];
        if (right_dst) right_dst[j] = right_src[j];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned long long e = *epoch_base + epoch;
        if (left_flag) st_release_sys(left_flag, e);
        if (right_flag) st_release_sys(right_flag, e);
    }
}

// Wait for the neighbours' epoch, then copy their payload from this rank's slots into the halo zones.  One CTA.
template <typename T>
__global__ void __launch_bounds__(256) halo_wait_kernel(const T* from_left_slot, const T* from_right_slot, long long count,
                                                        T* __restrict__ left_halo, T* __restrict__ right_halo,
                                                        const unsigned long long* flag_left, const unsigned long long* flag_right,
                                                        const unsigned long long* epoch_base, unsigned long long epoch, int* err) {
    __shared__ int ok;
    if (threadIdx.x == 0) {
        const unsigned long long e = *epoch_base + epoch;
        bool good = true;
        if (flag_left) good = peer_wait(flag_left, e) && good;
        if (flag_right) good = peer_wait(flag_right, e) && good;
        if (!good && err) atomicExch(err, 1);
        ok = good ? 1 : 0;
    }
    __syncthreads();
    if (!ok) return;
    for (long long j = threadIdx.x; j < count; j += blockDim.x) {
        if (from_left_slot) left_halo[j] = __ldcg(from_left_slot + j);
        if (from_right_slot) right_halo[j] = __ldcg(from_right_slot + j);
    }
}

// Push and wait in one launch (real ranks only: every rank's kernel is resident at the same time; ranks emulated on
// one stream must push first and wait afterwards, or the first one would wait for a kernel queued behind it).
template <typename T>
__global__ void __launch_bounds__(256) halo_exchange_kernel(const T* __restrict__ left_src, const T* __restrict__ right_src,
                                                            long long count, T* left_dst, T* right_dst,
                                                            unsigned long long* left_flag, unsigned long long* right_flag,
                                                            const T* from_left_slot, const T* from_right_slot,
                                                            T* __restrict__ left_halo, T* __restrict__ right_halo,
                                                            const unsigned long long* flag_left, const unsigned long long* flag_right,
                                                            const unsigned long long* epoch_base, unsigned long long epoch, int* err) {
    __shared__ int ok;
    for (long long j = threadIdx.x; j < count; j += blockDim.x) {
        if (left_dst) left_dst[j] = left_src[j];
        if (right_dst) right_dst[j] = right_src[j];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned long long e = *epoch_base + epoch;
        if (left_flag) st_release_sys(left_flag, e);
        if (right_flag) st_release_sys(right_flag, e);
        bool good = true;
        if (flag_left) good = peer_wait(flag_left, e) && good;
        if (flag_right) good = peer_wait(flag_right, e) && good;
        if (!good && err) atomicExch(err, 1);
        ok = good ? 1 : 0;
    }
    __syncthreads();
    if (!ok) return;
    for (long long j = threadIdx.x; j < count; j += blockDim.x) {
        if (from_left_slot) left_halo[j] = __ldcg(from_left_slot + j);
        if (from_right_slot) right_halo[j] = __ldcg(from_right_slot + j);
    }
}

__global__ void epoch_advance_kernel(unsigned long long* epoch_base, unsigned long long by) { *epoch_base += by; }

// All-reduce (sum) of `rows` values over `world` ranks in a fixed order (rank 0 first): bit-identical on every rank.
// slots[r] -> rank r's slot array [2][world][rows] (parity, source rank, row); flags[r] -> rank r's flags [world].
template <typename T>
__global__ void __launch_bounds__(256) peer_sum_kernel(const T* __restrict__ local, int rows, int rank, int world,
                                                       T* const* __restrict__ slots, unsigned long long* const* __restrict__ flags,
                                                       const unsigned long long* epoch_base, unsigned long long epoch,
                                                       T* __restrict__ out, int* err) {
    __shared__ int ok;
    const unsigned long long e = *epoch_base + epoch;
    const long long par = (long long)(e & 1ull) * world * rows;
    for (int r = 0; r < world; ++r) {
        T* dst = slots[r] + par + (long long)rank * rows;
        for (int j = threadIdx.x; j < rows; j += blockDim.x) dst[j] = local[j];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < world) st_release_sys(flags[threadIdx.x] + rank, e);
    if (threadIdx.x == 0) ok = 1;
    __syncthreads();
    if (threadIdx.x < world) {
        if (!peer_wait(flags[rank] + threadIdx.x, e)) {
            ok = 0;
            if (err) atomicExch(err, 1);
        }
    }
    __syncthreads();
    if (!ok) return;
    const T* mine = slots[rank] + par;
    for (int j = threadIdx.x; j < rows; j += blockDim.x) {
        T acc = __ldcg(mine + j);
        for (int r = 1; r < world; ++r) acc += __ldcg(mine + (long long)r * rows + j);
        out[j] = acc;
    }
}

}  // namespace oz

using namespace oz;

extern "C" {

OZ_API int oz_peer_halo_push(int dtype, const void* left_src, const void* right_src, int64_t count, void* left_dst,
                             void* right_dst, uint64_t* left_flag, uint64_t* right_flag, const uint64_t* epoch_base,
                             uint64_t epoch, void* stream) {
    OZ_REQUIRE(dtype == OZ_F64 || dtype == OZ_F32, "unknown dtype %d", dtype);
    OZ_REQUIRE(count >= 0 && epoch_base, "halo push: bad arguments");
    OZ_REQUIRE((!left_dst || (left_src && left_flag)) && (!right_dst || (right_src && right_flag)),
               "halo push: a destination needs its source and flag");
    cudaStream_t st = (cudaStream_t)stream;
    auto lf = (unsigned long long*)left_flag, rf = (unsigned long long*)right_flag;
    auto eb = (const unsigned long long*)epoch_base;
    if (dtype == OZ_F64)
        halo_push_kernel<double><<<1, 256, 0, st>>>((const double*)left_src, (const double*)right_src, count, (double*)left_dst,
                                                   (double*)right_dst, lf, rf, eb, epoch);
    else
        halo_push_kernel<float><<<1, 256, 0, st>>>((const float*)left_src, (const float*)right_src, count, (float*)left_dst,
                                                  (float*)right_dst, lf, rf, eb, epoch);
    OZ_CUDA_CHECK(cudaGetLastError());
    return 0;
}

OZ_API int oz_peer_halo_wait(int dtype, const void* from_left_slot, const void* from_right_slot, int64_t count,
                             void* left_halo, void* right_halo, const uint64_t* flag_left, const uint64_t* flag_right,
                             const uint64_t* epoch_base, uint64_t epoch, int32_t* err, void* stream) {
    OZ_REQUIRE(dtype == OZ_F64 || dtype == OZ_F32, "unknown dtype %d", dtype);
    OZ_REQUIRE(count >= 0 && epoch_base, "halo wait: bad arguments");
    OZ_REQUIRE((!from_left_slot || (left_halo && flag_left)) && (!from_right_slot || (right_halo && flag_right)),
               "halo wait: a slot needs its destination and flag");
    cudaStream_t st = (cudaStream_t)stream;
    auto fl = (const unsigned long long*)flag_left, fr = (const unsigned long long*)flag_right;
    auto eb = (const unsigned long long*)epoch_base;
    if (dtype == OZ_F64)
        halo_wait_kernel<double><<<1, 256, 0, st>>>((const double*)from_left_slot, (const double*)from_right_slot, count,
                                                   (double*)left_halo, (double*)right_halo, fl, fr, eb, epoch, err);
    else
        halo_wait_kernel<float><<<1, 256, 0, st>>>((const float*)from_left_slot, (const float*)from_right_slot, count,
                                                  (float*)left_halo, (float*)right_halo, fl, fr, eb, epoch, err);
    OZ_CUDA_CHECK(cudaGetLastError());
    return 0;
}

OZ_API int oz_peer_halo_exchange(int dtype, const void* left_src, const void* right_src, int64_t count, void* left_dst,
                                 void* right_dst, uint64_t* left_flag, uint64_t* right_flag, const void* from_left_slot,
                                 const void* from_right_slot, void* left_halo, void* right_halo, const uint64_t* flag_left,
                                 const uint64_t* flag_right, const uint64_t* epoch_base, uint64_t epoch, int32_t* err,
                                 void* stream) {
    OZ_REQUIRE(dtype == OZ_F64 || dtype == OZ_F32, "unknown dtype %d", dtype);
    OZ_REQUIRE(count >= 0 && epoch_base, "halo exchange: bad arguments");
    OZ_REQUIRE((!left_dst || (left_src && left_flag && from_left_slot && left_halo && flag_left)) &&
                   (!right_dst || (right_src && right_flag && from_right_slot && right_halo && flag_right)),
               "halo exchange: a neighbour needs source, slot, halo and both flags");
    cudaStream_t st = (cudaStream_t)stream;
    auto lf = (unsigned long long*)left_flag, rf = (unsigned long long*)right_flag;
    auto fl = (const unsigned long long*)flag_left, fr = (const unsigned long long*)flag_right;
    auto eb = (const unsigned long long*)epoch_base;
    if (dtype == OZ_F64)
        halo_exchange_kernel<double><<<1, 256, 0, st>>>((const double*)left_src, (const double*)right_src, count, (double*)left_dst,
                                                       (double*)right_dst, lf, rf, (const double*)from_left_slot,
                                                       (const double*)from_right_slot, (double*)left_halo, (double*)right_halo, fl, fr,
                                                       eb, epoch, err);
    else
        halo_exchange_kernel<float><<<1, 256, 0, st>>>((const float*)left_src, (const float*)right_src, count, (float*)left_dst,
                                                      (float*)right_dst, lf, rf, (const float*)from_left_slot,
                                                      (const float*)from_right_slot, (float*)left_halo, (float*)right_halo, fl, fr, eb,
                                                      epoch, err);
    OZ_CUDA_CHECK(cudaGetLastError());
    return 0;
}

OZ_API int oz_peer_epoch_advance(uint64_t* epoch_base, uint64_t by, void* stream) {
    OZ_REQUIRE(epoch_base != nullptr, "null epoch base");
    epoch_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((unsigned long long*)epoch_base, by);
    OZ_CUDA_CHECK(cudaGetLastError());
    return 0;
}

OZ_API int oz_peer_sum(int dtype, const void* local, int32_t rows, int32_t rank, int32_t world, void* const* slots,
                       uint64_t* const* flags, const uint64_t* epoch_base, uint64_t epoch, void* out, int32_t* err,
                       void* stream) {
    OZ_REQUIRE(dtype == OZ_F64 || dtype == OZ_F32, "unknown dtype %d", dtype);
    OZ_REQUIRE(local && out && slots && flags && epoch_base, "peer sum: null argument");
    OZ_REQUIRE(rows >= 1 && world >= 1 && world <= 64 && rank >= 0 && rank < world, "peer sum: bad sizes");
    cudaStream_t st = (cudaStream_t)stream;
    auto fl = (unsigned long long* const*)flags;
    auto eb = (const unsigned long long*)epoch_base;
    if (dtype == OZ_F64)
        peer_sum_kernel<double><<<1, 256, 0, st>>>((const double*)local, rows, rank, world, (double* const*)slots, fl, eb, epoch,
                                                  (double*)out, err);
    else
        peer_sum_kernel<float><<<1, 256, 0, st>>>((const float*)local, rows, rank, world, (float* const*)slots, fl, eb, epoch,
                                                 (float*)out, err);
    OZ_CUDA_CHECK(cudaGetLastError());
    return 0;
}

}  // extern "C"
