// Collocation: the GLM equations as equality constraints on design variables, and (dc)^T mu.
//
// Reference: ozone/integrators/collocation.py:33-57, ozone/collections.py:196-232.
//   Ydv [n][s][d][N]  stage unknowns,  ydv [n][d][N] state unknowns y_1..y_n,  yfull = [y0; ydv]
//   c1 = A hF + U yfull - Ydv                       [n][s][d][N]
//   c2 = (B hF + V yfull - yfull)[d:]               [n][d][N]    = h sum_i b_i F_ki + y_k - y_{k+1}
// Everything is independent across (trajectory, step): the residual kernel runs one thread per
// (trajectory, step); the VJP kernel gives each thread a short run of consecutive steps so that the
// lerp-transpose of the parameter gradient (p_bar[k] gets (1-c) parts of step k and c parts of step
// k-1) is assembled by one owner per entry -- no atomics, deterministic.
#pragma once
#include <cstdlib>
#include "oz_common.cuh"
#include "oz_tm_kernels.cuh"

namespace oz {

template <typename T>
struct ColArgs {
    Tableau<T> tab;
    long long N;
    int n, chunk, nchunks;
    int h_shared, p_shared, th_shared;
    const T *y0, *h, *p, *th;
    const T *Y, *y;
    T *c1, *c2;
    const T *mu1, *mu2;
    T *Y_bar, *y_bar, *y0_bar, *h_bar, *p_bar, *th_bar;
    T *ws_h, *ws_p, *ws_th;
    long long W;  // warps along the trajectory axis
};

template <class R, typename T>
OZ_D void col_load_p(const ColArgs<T>& a, int k, long long i, T* p) {
#pragma unroll
    for (int c = 0; c < R::P; ++c) p[c] = ld_in(a.p, (long long)k * R::P + c, a.N, i, a.p_shared);
}

#ifndef OZ_COLLOC_CHUNK   /* consecutive steps per thread of the VJP kernel (each run recomputes the step before it for p_bar) */
#define OZ_COLLOC_CHUNK 10
#endif
// resident blocks the residual kernel is compiled for: its live state is the S x D accumulators of c1 (lunar ascent, measured:
// RK4 S D = 28 -> 3 blocks 0.180 ms / 4 blocks 0.188 ms; GL6 S D = 21 -> 0.146 / 0.134 ms)
template <class R, int S>
constexpr int colloc_res_blocks() {
    return S * R::D <= 21 ? 4 : 3;
}
#ifndef OZ_MB_COLLOC
#define OZ_MB_COLLOC 3
#endif
template <class R, typename T, int S>
__global__ void __launch_bounds__(128, (colloc_res_blocks<R, S>())) colloc_residual_kernel(const ColArgs<T> a) {
    using DM = Dims<R>;
    constexpr int D = DM::D;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int k = blockIdx.y;
    if (i >= a.N) return;
    const long long N = a.N;
    const Tableau<T>& tab = a.tab;
    // c1 / c2 are accumulated stage by stage: once F_j is known it is added to every c1_s (weight h a_sj) and to c2
    // (weight h b_j), so only ONE stage's Y and F are live at a time instead of all S of them
    T th[DM::S1], plo[DM::P1], phi[DM::P1], c1[S][D], c2[D];
#pragma unroll
    for (int c = 0; c < DM::SP; ++c) th[c] = ld_in(a.th, c, N, i, a.th_shared);
    col_load_p<R>(a, k, i, plo);
    col_load_p<R>(a, k + 1, i, phi);
    const T h = ld_in(a.h, k, N, i, a.h_shared);
#pragma unroll
    for (int c = 0; c < D; ++c) {
        const T yk = (k == 0) ? __ldg(a.y0 + c * N + i) : __ldg(a.y + ((long long)(k - 1) * D + c) * N + i);
        c2[c] = yk - __ldg(a.y + ((long long)k * D + c) * N + i);
#pragma unroll
        for (int s = 0; s < S; ++s) c1[s][c] = yk;
    }
#pragma unroll
    for (int j = 0; j < S; ++j) {
        T Yj[D], Fj[D], pj[DM::P1];
#pragma unroll
        for (int c = 0; c < DM::P; ++c) pj[c] = plo[c] + tab.C[j] * (phi[c] - plo[c]);
#pragma unroll
        for (int c = 0; c < D; ++c) {
            Yj[c] = __ldg(a.Y + (((long long)k * S + j) * D + c) * N + i);
            c1[j][c] -= Yj[c];
        }
        R::f(Yj, pj, th, Fj);
#pragma unroll
        for (int c = 0; c < D; ++c) {
            const T hf = h * Fj[c];
            c2[c] += tab.B[j] * hf;
#pragma unroll
            for (int s = 0; s < S; ++s) c1[s][c] += tab.a(s, j) * hf;
        }
    }
#pragma unroll
    for (int c = 0; c < D; ++c) {
#pragma unroll
        for (int s = 0; s < S; ++s) a.c1[(((long long)k * S + s) * D + c) * N + i] = c1[s][c];
        a.c2[((long long)k * D + c) * N + i] = c2[c];
    }
}

// per-stage cotangent of hF:  W_j = sum_i a_ij mu1_ki + b_j mu2_k ;  G_j = h W_j
template <class R, typename T, int S>
OZ_D void col_step_vjp(const ColArgs<T>& a, int k, long long i, const T* th, const T* plo, const T* phi, T h,
                       bool write_Y, T& hb, T* pb_lo, T* pb_hi, T* tb, T* mu2_prev = nullptr) {
    using DM = Dims<R>;
    constexpr int D = DM::D;
    const long long N = a.N;
    const Tableau<T>& tab = a.tab;
    T ps[S][DM::P1], mu1[S][D], mu2[D];
    stage_params<R, T, S>(tab, plo, phi, ps);
#pragma unroll
    for (int c = 0; c < D; ++c) {
        mu2[c] = __ldg(a.mu2 + ((long long)k * D + c) * N + i);
#pragma unroll
        for (int s = 0; s < S; ++s) mu1[s][c] = __ldg(a.mu1 + (((long long)k * S + s) * D + c) * N + i);
    }
    // yfull[k] enters c1_k (+, every stage), c2_k (+) and c2_{k-1} (-): its gradient is formed here, while this step's
    // multiplier rows are in registers (mu2_prev carries mu2 of step k - 1 from call to call)
    if (mu2_prev) {
#pragma unroll
        for (int c = 0; c < D; ++c) {
            T g = mu2[c];
#pragma unroll
            for (int s = 0; s < S; ++s) g += mu1[s][c];
            if (write_Y) {
                if (k == 0) {
                    if (a.y0_bar) a.y0_bar[c * N + i] = g;
                } else if (a.y_bar) {
                    a.y_bar[((long long)(k - 1) * D + c) * N + i] = g - mu2_prev[c];
                }
            }
            mu2_prev[c] = mu2[c];
        }
    }
    hb = T(0);
#pragma unroll
    for (int c = 0; c < DM::P1; ++c) {
        pb_lo[c] = T(0);
        pb_hi[c] = T(0);
    }
#pragma unroll
    for (int j = 0; j < S; ++j) {
        T Yj[D], Fj[D], G[D], yb[D], pbj[DM::P1];
#pragma unroll
        for (int c = 0; c < DM::P1; ++c) pbj[c] = T(0);
#pragma unroll
        for (int c = 0; c < D; ++c) {
            Yj[c] = __ldg(a.Y + (((long long)k * S + j) * D + c) * N + i);
            T w = tab.B[j] * mu2[c];
#pragma unroll
            for (int s = 0; s < S; ++s) w += tab.a(s, j) * mu1[s][c];
            G[c] = w;  // W_j for now
            yb[c] = -mu1[j][c];
        }
        R::f(Yj, ps[j], th, Fj);
#pragma unroll
        for (int c = 0; c < D; ++c) {
            hb += Fj[c] * G[c];
            G[c] *= h;
        }
        R::vjp(Yj, ps[j], th, G, yb, pbj, tb);
        if (write_Y && a.Y_bar)
#pragma unroll
            for (int c = 0; c < D; ++c) a.Y_bar[(((long long)k * S + j) * D + c) * N + i] = yb[c];
#pragma unroll
        for (int c = 0; c < DM::P; ++c) {
            pb_lo[c] += (T(1) - tab.C[j]) * pbj[c];
            pb_hi[c] += tab.C[j] * pbj[c];
        }
    }
}

template <class R, typename T, int S>
__global__ void __launch_bounds__(128, OZ_MB_COLLOC) colloc_vjp_kernel(const ColArgs<T> a) {
    using DM = Dims<R>;
    constexpr int D = DM::D;
    const long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = gi < a.N;
    const long long N = a.N, i = valid ? gi : N - 1;
    const int lane = threadIdx.x & 31;
    const long long wx = gi >> 5;
    const int chunk_id = blockIdx.y;
    const int k_lo = chunk_id * a.chunk, k_hi = min(a.n, k_lo + a.chunk);
    T th[DM::S1], tb[DM::S1], tb_dummy[DM::S1], plo[DM::P1], phi[DM::P1], pb_lo[DM::P1], pb_hi[DM::P1], carry[DM::P1];
#pragma unroll
    for (int c = 0; c < DM::SP; ++c) th[c] = ld_in(a.th, c, N, i, a.th_shared);
#pragma unroll
    for (int c = 0; c < DM::S1; ++c) tb[c] = T(0);
#pragma unroll
    for (int c = 0; c < DM::P1; ++c) carry[c] = T(0);

    auto emit_p = [&](int k, const T* v) {  // p_bar[k] complete
        if (!a.p_bar) return;
#pragma unroll
        for (int c = 0; c < DM::P; ++c) {
            if (a.p_shared) {
                const T s = warp_sum(valid ? v[c] : T(0));
                if (lane == 0) a.ws_p[((long long)k * DM::P + c) * a.W + wx] = s;
            } else if (valid) {
                a.p_bar[((long long)k * DM::P + c) * N + i] = v[c];
            }
        }
    };

    // hi-part of the step just before this run (owned by the previous run for everything else)
    if (k_lo > 0 && DM::P > 0) {
        T hb;
        col_load_p<R>(a, k_lo - 1, i, plo);
        col_load_p<R>(a, k_lo, i, phi);
        const T h = ld_in(a.h, k_lo - 1, N, i, a.h_shared);
        col_step_vjp<R, T, S>(a, k_lo - 1, i, th, plo, phi, h, false, hb, pb_lo, pb_hi, tb_dummy);
#pragma unroll
        for (int c = 0; c < DM::P; ++c) carry[c] = pb_hi[c];
    }
    col_load_p<R>(a, k_lo, i, plo);
    T mu2_prev[D];   // mu2 of step k - 1: subtracted from y_bar[k - 1] (c2_{k-1} holds -yfull[k])
#pragma unroll
    for (int c = 0; c < D; ++c) mu2_prev[c] = k_lo > 0 ? __ldg(a.mu2 + ((long long)(k_lo - 1) * D + c) * N + i) : T(0);
    for (int k = k_lo; k < k_hi; ++k) {
        T hb;
        col_load_p<R>(a, k + 1, i, phi);
        const T h = ld_in(a.h, k, N, i, a.h_shared);
        col_step_vjp<R, T, S>(a, k, i, th, plo, phi, h, valid, hb, pb_lo, pb_hi, tb, mu2_prev);
#pragma unroll
        for (int c = 0; c < DM::P; ++c) carry[c] += pb_lo[c];
        emit_p(k, carry);
#pragma unroll
        for (int c = 0; c < DM::P; ++c) {
            carry[c] = pb_hi[c];
            plo[c] = phi[c];
        }
        if (a.h_bar) {
            if (a.h_shared) {
                const T s = warp_sum(valid ? hb : T(0));
                if (lane == 0) a.ws_h[(long long)k * a.W + wx] = s;
            } else if (valid) {
                a.h_bar[(long long)k * N + i] = hb;
            }
        }
    }
    if (k_hi == a.n) {
        emit_p(a.n, carry);
        if (valid && a.y_bar)  // y_n enters only c2_{n-1} (-)
#pragma unroll
            for (int c = 0; c < D; ++c) a.y_bar[((long long)(a.n - 1) * D + c) * N + i] = -mu2_prev[c];
    }
    if (a.th_bar) {
#pragma unroll
        for (int c = 0; c < DM::SP; ++c) {
            if (a.th_shared) {
                const T s = warp_sum(valid ? tb[c] : T(0));
                if (lane == 0) a.ws_th[((long long)c * a.nchunks + chunk_id) * a.W + wx] = s;
            } else if (valid) {   // per-trajectory static parameter: one partial per (time chunk, trajectory)
                a.ws_th[((long long)c * a.nchunks + chunk_id) * N + i] = tb[c];
            }
        }
    }
}

// th_bar[c][i] = sum over the time chunks of one trajectory's partials (fixed order)
template <typename T>
__global__ void __launch_bounds__(256) sum_chunks_kernel(const T* __restrict__ part, int nchunks, long long N,
                                                         T* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const long long c = blockIdx.y;
    T acc = T(0);
    for (int q = 0; q < nchunks; ++q) acc += part[(c * nchunks + q) * N + i];
    out[c * N + i] = acc;
}

struct ColWs {
    long long W, rows_h, rows_p, rows_th, th_elems, off_h, off_p, off_th, total;
    int chunk, nchunks;
};

template <class R>
inline ColWs colloc_ws_layout(const oz_problem* pb) {
    ColWs w;
    const int block = 128;
    // runs of about OZ_COLLOC_CHUNK steps, balanced so that the last run is not a short one (n = 100: 10 runs of 10;
    // 13 runs of 8 with a ragged last one measured 3 % slower, profiles/r2_colloc_vjp_chunk_sweep.txt)
    int target = OZ_COLLOC_CHUNK;
    if (const char* e = getenv("OZ_COLLOC_CHUNK")) target = atoi(e) > 0 ? atoi(e) : OZ_COLLOC_CHUNK;   // experiments
    int runs = (pb->n + target / 2) / target;
    if (runs < 1) runs = 1;
    w.chunk = (pb->n + runs - 1) / runs;
    w.nchunks = (pb->n + w.chunk - 1) / w.chunk;
    w.W = ((pb->N + block - 1) / block) * (block / 32);
    w.rows_h = (pb->h_shared && pb->h_bar) ? pb->n : 0;
    w.rows_p = (pb->p_shared && pb->p_bar) ? (long long)(pb->n + 1) * R::P : 0;
    w.rows_th = pb->th_bar ? (long long)R::S : 0;  // always reduced over the time chunks
    // shared th: one partial per (chunk, warp); per-trajectory th: one per (chunk, trajectory)
    w.th_elems = w.rows_th * w.nchunks * (pb->th_shared ? w.W : (long long)pb->N);
    w.off_h = 0;
    w.off_p = w.off_h + w.rows_h * w.W;
    w.off_th = w.off_p + w.rows_p * w.W;
    w.total = w.off_th + w.th_elems;
    return w;
}

template <class R, typename T, int S>
int run_colloc(const oz_problem* pb, int mode, size_t* ws_bytes, const void* Y, const void* y, const void* mu1,
               const void* mu2, void* o1, void* o2) {
    const ColWs w = colloc_ws_layout<R>(pb);
    if (mode == 2) {
        *ws_bytes = (size_t)w.total * sizeof(T);
        return 0;
    }
    ColArgs<T> a;
    a.tab = make_tableau<T>(pb);
    a.N = pb->N;
    a.n = pb->n;
    a.chunk = w.chunk;
    a.nchunks = w.nchunks;
    a.h_shared = pb->h_shared;
    a.p_shared = pb->p_shared;
    a.th_shared = pb->th_shared;
    a.y0 = (const T*)pb->y0;
    a.h = (const T*)pb->h;
    a.p = (const T*)pb->p;
    a.th = (const T*)pb->th;
    a.Y = (const T*)Y;
    a.y = (const T*)y;
    a.W = w.W;
    OZ_REQUIRE(Y && y, "stage and state design variables are required");
    OZ_REQUIRE(pb->n <= 65535, "at most 65535 steps per collocation launch");
    cudaStream_t stream = (cudaStream_t)pb->stream;
    const unsigned gx = (unsigned)((pb->N + 127) / 128);
    if (mode == 0) {
        OZ_REQUIRE(o1 && o2, "c1 and c2 outputs are required");
        a.c1 = (T*)o1;
        a.c2 = (T*)o2;
        colloc_residual_kernel<R, T, S><<<dim3(gx, pb->n), 128, 0, stream>>>(a);
        OZ_CUDA_CHECK(cudaGetLastError());
        return 0;
    }
    OZ_REQUIRE(mu1 && mu2, "multipliers mu1, mu2 are required");
    OZ_REQUIRE((size_t)w.total * sizeof(T) <= pb->ws_bytes && (w.total == 0 || pb->ws),
               "collocation VJP workspace too small: need %zu bytes, got %zu", (size_t)w.total * sizeof(T), pb->ws_bytes);
    a.mu1 = (const T*)mu1;
    a.mu2 = (const T*)mu2;
    a.Y_bar = (T*)o1;
    a.y_bar = (T*)o2;
    a.y0_bar = (T*)pb->y0_bar;
    a.h_bar = (T*)pb->h_bar;
    a.p_bar = (T*)pb->p_bar;
    a.th_bar = (T*)pb->th_bar;
    T* ws = (T*)pb->ws;
    a.ws_h = ws + w.off_h;
    a.ws_p = ws + w.off_p;
    a.ws_th = ws + w.off_th;
    colloc_vjp_kernel<R, T, S><<<dim3(gx, w.nchunks), 128, 0, stream>>>(a);
    OZ_CUDA_CHECK(cudaGetLastError());
    if (w.rows_h) reduce_rows_kernel<T><<<(unsigned)w.rows_h, 256, 0, stream>>>(a.ws_h, w.W, a.h_bar);
    if (w.rows_p) reduce_rows_kernel<T><<<(unsigned)w.rows_p, 256, 0, stream>>>(a.ws_p, w.W, a.p_bar);
    if (w.rows_th && pb->th_shared)
        reduce_rows_kernel<T><<<(unsigned)w.rows_th, 256, 0, stream>>>(a.ws_th, w.W * w.nchunks, a.th_bar);
    if (w.rows_th && !pb->th_shared)
        sum_chunks_kernel<T><<<dim3((unsigned)((pb->N + 255) / 256), (unsigned)w.rows_th), 256, 0, stream>>>(a.ws_th, w.nchunks,
                                                                                                           pb->N, a.th_bar);
    OZ_CUDA_CHECK(cudaGetLastError());
    return 0;
}

template <class R>
int dispatch_colloc(const oz_problem* pb, int mode, size_t* ws_bytes, const void* Y, const void* y, const void* mu1,
                    const void* mu2, void* o1, void* o2) {
#define OZ_COL_CASE(SS)                                                                                          \
    if constexpr (OZ_WANT_S(SS) && OZ_WANT_KIND(2)) {                                                            \
        if (pb->num_stages == SS) {                                                                              \
            if constexpr (OZ_WANT_F64(true))                                                                     \
                if (pb->dtype == OZ_F64) return run_colloc<R, double, SS>(pb, mode, ws_bytes, Y, y, mu1, mu2, o1, o2); \
            if constexpr (OZ_WANT_F64(false))                                                                    \
                if (pb->dtype == OZ_F32) return run_colloc<R, float, SS>(pb, mode, ws_bytes, Y, y, mu1, mu2, o1, o2);  \
        }                                                                                                        \
    }
    OZ_COL_CASE(1)
    OZ_COL_CASE(2)
    OZ_COL_CASE(3)
    OZ_COL_CASE(4)
    OZ_COL_CASE(7)
#undef OZ_COL_CASE
    set_error("no collocation kernel for %d stages / dtype %d in this build", pb->num_stages, pb->dtype);
    return -1;
}

}  // namespace oz
