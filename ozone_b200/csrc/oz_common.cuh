// Shared device/host helpers for the ozone_b200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include "../../include/ozone_b200.h"

#define OZ_HD __host__ __device__ __forceinline__
#define OZ_D __device__ __forceinline__

// A plugin build instantiates one combination only: -DOZ_ONLY_S=4 -DOZ_ONLY_IMPL=0 -DOZ_ONLY_F64=1
#ifdef OZ_ONLY_S
#define OZ_WANT_S(s) ((s) == OZ_ONLY_S)
#else
#define OZ_WANT_S(s) true
#endif
#ifdef OZ_ONLY_IMPL
#define OZ_WANT_IMPL(b) ((b) == (OZ_ONLY_IMPL != 0))
#else
#define OZ_WANT_IMPL(b) true
#endif
#ifdef OZ_ONLY_F64
#define OZ_WANT_F64(b) ((b) == (OZ_ONLY_F64 != 0))
#else
#define OZ_WANT_F64(b) true
#endif
#ifdef OZ_ONLY_KIND      /* 0 time marching, 1 picard, 2 collocation */
#define OZ_WANT_KIND(k) ((k) == OZ_ONLY_KIND)
#else
#define OZ_WANT_KIND(k) true
#endif

namespace oz {

// Thread-local error text behind oz_last_error().
void set_error(const char* fmt, ...);

#define OZ_CUDA_CHECK(expr)                                                          \
    do {                                                                             \
        cudaError_t _e = (expr);                                                     \
        if (_e != cudaSuccess) {                                                     \
            oz::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return -2;                                                               \
        }                                                                            \
    } while (0)

#define OZ_REQUIRE(cond, ...)                   \
    do {                                        \
        if (!(cond)) {                          \
            oz::set_error(__VA_ARGS__);         \
            return -1;                          \
        }                                       \
    } while (0)

constexpr int kMaxS = OZ_MAX_STAGES;
constexpr size_t kSmemBudget = 200 * 1024;  // dynamic shared memory we size blocks for (227 KB is the hard cap)

// Tableau as seen by a kernel: passed BY VALUE inside the kernel parameter block, i.e. it sits in
// the constant bank and every access is a uniform (warp-broadcast) load.  No __constant__ symbol,
// so concurrent solves with different methods on different streams cannot race.
template <typename T>
struct Tableau {
    T A[kMaxS * kMaxS];
    T B[kMaxS];
    T C[kMaxS];
    OZ_HD T a(int i, int j) const { return A[i * kMaxS + j]; }
};

template <typename T>
inline Tableau<T> make_tableau(const oz_problem* pb) {
    Tableau<T> t;
    for (int i = 0; i < kMaxS * kMaxS; ++i) t.A[i] = (T)pb->A[i];
    for (int i = 0; i < kMaxS; ++i) {
        t.B[i] = (T)pb->B[i];
        t.C[i] = (T)pb->C[i];
    }
    return t;
}

// Everything a time-marching kernel needs, typed.
template <typename T>
struct TmArgs {
    Tableau<T> tab;
    long long N;
    int n;
    int num_inner;  // 0: full history available; > 0: checkpoints every num_inner steps
    int h_shared, p_shared, th_shared;
    int chain;      // explicit tableau whose only nonzero A entries are a(i, i-1) (RK4, midpoint, Euler): one FMA per stage
    T newton_tol;
    int newton_max_iter;
    const T* y0;
    const T* h;
    const T* p;
    const T* th;
    T* hist;
    T* ckpt;
    T* yfinal;
    T* prof;
    T* field;
    int* status;
    int* iters;
    const T* hist_bar;
    const T* final_bar;
    const T* prof_bar;
    const T* field_bar;
    T* y0_bar;
    T* h_bar;
    T* p_bar;
    T* th_bar;
    T* stages;    // optional [n][s][d][N] cache of converged implicit stage values (forward writes, adjoint reads)
    T* ws;        // adjoint scratch (see tm_adjoint_ws_layout)
    long long W;  // number of warps in the grid = partial sums per reduced quantity
};

template <typename T>
inline TmArgs<T> make_tm_args(const oz_problem* pb) {
    TmArgs<T> a;
    a.tab = make_tableau<T>(pb);
    a.N = pb->N;
    a.n = pb->n;
    a.num_inner = pb->num_inner;
    a.chain = pb->is_explicit ? 1 : 0;
    for (int i = 0; i < pb->num_stages; ++i)
        for (int j = 0; j < pb->num_stages; ++j)
            if (j != i - 1 && pb->A[i * kMaxS + j] != 0.0) a.chain = 0;
    a.h_shared = pb->h_shared;
    a.p_shared = pb->p_shared;
    a.th_shared = pb->th_shared;
    a.newton_tol = (T)pb->newton_tol;
    a.newton_max_iter = pb->newton_max_iter;
    a.y0 = (const T*)pb->y0;
    a.h = (const T*)pb->h;
    a.p = (const T*)pb->p;
    a.th = (const T*)pb->th;
    a.hist = (T*)pb->hist;
    a.ckpt = (T*)pb->ckpt;
    a.yfinal = (T*)pb->yfinal;
    a.prof = (T*)pb->prof;
    a.field = (T*)pb->field;
    a.status = pb->status;
    a.iters = pb->iters;
    a.hist_bar = (const T*)pb->hist_bar;
    a.final_bar = (const T*)pb->final_bar;
    a.prof_bar = (const T*)pb->prof_bar;
    a.field_bar = (const T*)pb->field_bar;
    a.y0_bar = (T*)pb->y0_bar;
    a.h_bar = (T*)pb->h_bar;
    a.p_bar = (T*)pb->p_bar;
    a.th_bar = (T*)pb->th_bar;
    a.stages = (T*)pb->stages;
    a.ws = (T*)pb->ws;
    a.W = 0;
    return a;
}

template <typename T>
OZ_D T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int N>
struct AtLeast1 {
    static constexpr int value = N > 0 ? N : 1;
};

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

int sm_count();

}  // namespace oz
