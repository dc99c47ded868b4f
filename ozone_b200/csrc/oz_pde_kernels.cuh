// Grid right-hand sides under TimeMarching / TimeMarchingCheckpoints with any EXPLICIT tableau: ONE trajectory whose
// d = nx state components are the cells of a 1-D grid coupled by a three-point stencil with zero ghost cells, so the
// parallel axis is SPACE.  The kernels are written against a "grid model" M (contract below, at PdeArgs): the built-in
// diffusion-reaction system of ozone/paper_examples/pde_control/pde_control.py:11-81 (oz_pde.cu), or the functor that
// ozone_b200/trace_grid.py generates from a Python callback written like that file (oz_grid_plugin_impl.cuh).
//
// One launch advances the whole grid by several steps with all s stages of each fused: a CTA loads its tile plus
// a halo of s cells per side and step, the stage values circulate through one shared-memory line (each
// stencil application invalidates one more halo cell per side), F_i and the state stay in registers, and
// only the interior of the last state (or of every state, when the history is kept) is written.  HBM /
// L2 traffic per cell-step is at most one read + one write (2w), not 2w per stage.
// The reverse launch re-derives the stages of each step the same way and then runs the transposed stage chain
// (the transposed stencil has the same access pattern on the multipliers) with a halo of 2s per step.  Grid-wide
// sums (field output, u_bar, h_bar) are per-CTA partials per step, reduced in a fixed order.
//
// `x_off`/`own_lo`/`own_hi` let a rank run the same kernels on a slab [x_off, x_off + nx) of a larger
// grid: reductions only count the cells it owns (see ozone_b200/parallel.py for the halo exchange).
#pragma once
#include "oz_common.cuh"
#include "oz_tm_launch.cuh"
#include <cstdlib>
#include "oz_linalg.cuh"

namespace oz {

#ifndef OZ_PDE_BLOCK
#define OZ_PDE_BLOCK 256
#endif
#ifndef OZ_PDE_CPT
#define OZ_PDE_CPT 4
#endif
#ifndef OZ_MB_PDE_ADJ
#define OZ_MB_PDE_ADJ 2
#endif
#ifndef OZ_PDE_FUSE   /* forward steps per launch (keep engine.GRID_STEPS_PER_LAUNCH in step) */
#define OZ_PDE_FUSE 4
#endif
constexpr int kPdeBlock = OZ_PDE_BLOCK;
constexpr int kPdeCpt = OZ_PDE_CPT;                // cells per thread
constexpr int kPdeExt = kPdeBlock * kPdeCpt;      // cells per CTA including halos
constexpr int kPdeLine = kPdeExt + 2;             // one shared-memory line: the tile plus a zero cell at either end
constexpr int kPdeFuse = OZ_PDE_FUSE;
constexpr int kPdeMaxP = 4;                       // controls of a grid model
#ifndef OZ_PDE_FUSE_ADJ   /* reverse steps per launch */
#define OZ_PDE_FUSE_ADJ 2
#endif
constexpr int kPdeFuseAdj = OZ_PDE_FUSE_ADJ;
#ifndef OZ_PDE_TINY_CELLS        /* grids up to this many cells run 3 cells per thread instead of 4 */
#define OZ_PDE_TINY_CELLS 160000
#endif
#ifndef OZ_PDE_SMALL_CELLS      /* grids up to this many cells (about one wave of CTAs) fuse more steps per launch */
#define OZ_PDE_SMALL_CELLS 600000
#endif
#ifndef OZ_PDE_FUSE_SMALL
#define OZ_PDE_FUSE_SMALL 16
#endif
#ifndef OZ_PDE_FUSE_ADJ_SMALL
#define OZ_PDE_FUSE_ADJ_SMALL 8
#endif

// ---- the grid model ---------------------------------------------------------------------------------------------
// Everything below is written against a model M: the built-in diffusion-reaction system (DiffusionReaction in oz_pde.cu)
// or the functor ozone_b200/trace_grid.py generates from a Python callback.  A model describes ONE cell of a
// three-point stencil with zero ghost cells whose neighbour weights do not depend on the state (constants, or coefficient
// fields; other boundary conditions are coefficient fields at the boundary cells, see trace_grid.py):
//   using T;                                        arithmetic type
//   static constexpr int NCOEF;                     per-cell coefficient fields read from HBM (a.coef, [NCOEF][nx_global])
//   static constexpr bool WINDOW;                   the control acts on the index window [h_lo, h_hi) (members of M)
//   static constexpr int P;                         scalar, batch-shared dynamic parameters (controls), 0 .. kPdeMaxP
//   T stage_u(ulo, du, c)                           one control at a stage abscissa (CTA-uniform)
//   template <bool WITH_DIAG> T f(xl, x, xr, mask, coef, us, etab, dg)    x_j' and (WITH_DIAG) dg = what jt() needs; us[P]
//   T jt(gl, g, gr, dg, coef)                       (J^T G)_j from G_{j-1}, G_j, G_{j+1}
//   T du(mask, coef, q), T du_scale()               d x_j' / d u_q = du * du_scale (must not depend on the state)
//   T q(y, coef), T dq(y, coef)                     integrand of the field output and its derivative
//   T ucost(own, u), T ducost(fb, u, q)             the controls' own term in the field output (u[P]) and its derivative
template <class M>
struct PdeArgs {
    using T = typename M::T;
    Tableau<T> tab;
    M rhs;
    const T* coef;       // [NCOEF][coef_stride] coefficient fields, indexed by the GLOBAL cell
    long long coef_stride;
    T field_scale;       // field output = field_scale * sum_cells q + sum_k ucost(u_k)
    long long nx;        // cells in this (local) domain; outside is zero
    long long x_off;     // global index of local cell 0 (for the control mask)
    long long own_lo, own_hi;  // local range whose cells count in reductions
    long long nx_global;
    int tile;            // interior cells per CTA (<= kPdeExt - 2*HALO), balanced over the resident CTA slots
    int halo;            // cells on either side of the interior (forward: >= nsub * s, reverse: >= 2 * nsub * s)
    int nsub;            // steps fused in this launch
    const T* y_in;
    T* y_out;            // state after the last fused step (or nullptr when rows_out takes every step)
    T* rows_out;         // optional: state after fused step j goes to rows_out + j * row_stride
    long long row_stride;
    const T* h;          // &h[k]: fused step j uses h[j]
    const T* u_lo;       // &p[k][0], &p[k+1][0] (rows of P controls): fused step j uses u_lo[j * P + q], u_hi[j * P + q]
    const T* u_hi;
    T* part_sq;          // partial sums of the new state squared over owned cells, row j (of sq_stride) per fused step, or nullptr
    long long sq_stride;
    // adjoint
    // reverse launch: steps k, k-1, .., k-nsub+1; fused step j reads y_in - j*row_stride, h[-j], u_lo[-j], u_hi[-j],
    // hist_bar_k - j*nx and writes its partial sums to row -j (of adj_stride) of part_h / part_ulo / part_uhi
    const T* lam_in;     // lambda_{k+1}
    T* lam_out;          // lambda_{k-nsub+1}
    const T* hist_bar_k; // direct cotangent of y_k or nullptr
    long long adj_stride;
    const T* field_bar;  // device scalar (or nullptr): cotangent of the field output
    T fb_scale;          // d field / d y_j = field_bar * fb_scale * dq(y_j)
    T* part_h;           // [ncta]
    T* part_ulo;         // [P][part_pstride]: control q's partials of this launch start at q * part_pstride
    T* part_uhi;
    long long part_pstride;
};

template <typename T>
OZ_D T block_sum(T v, T* sm) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sm[w] = v;
    __syncthreads();
    T r = T(0);
    if (threadIdx.x < 32) {
        r = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : T(0);
        r = warp_sum(r);
    }
    return r;  // valid in warp 0
}

// ---- exp and reciprocal of the reaction term ------------------------------------------------------------------
// The fused step evaluates exp(-gamma / (1 + x)) once per cell and stage.  The library exp costs ~20 FP64
// instructions plus ~30 uniform-register moves for its 64-bit constants (they do not fit in the uniform file next
// to the tableau, so every inlined copy re-creates them: 24 % of all issued instructions were UMOV), and the IEEE
// division another ~10 with a slow-path branch.  Here:
//   exp(t) = 2^k 2^(j/32) e^r,  round(32 t log2 e) = 32 k + j,  |r| <= ln2 / 64,  e^r by a degree-6 Taylor sum
// (truncation 3e-18 relative), 2^(j/32) from a 32-entry shared-memory table: 11 FP64 instructions, 9 constants.
// 1 / (1 + x) is the reciprocal seed plus two Newton steps (oz_linalg.cuh).  Relative error of both ~2e-16.
constexpr int kExpTab = 32;

template <typename T>
OZ_D void pde_exp_table(T* tab) {  // caller synchronises before the first pde_exp; fp32 does not use the table
    if constexpr (sizeof(T) == 8) {
        if (threadIdx.x < kExpTab) tab[threadIdx.x] = (T)exp2((double)threadIdx.x / kExpTab);
    }
}

// fp32: the hardware exponential (ex2.approx after one multiply, relative error ~4e-7 for the arguments met here,
// -gamma/(1+x) in [-2, -0.4]) and the approximate reciprocal (1 ulp): both far inside the 1e-5 the fp32 path promises,
// and ~4 instructions instead of ~30 for expf + an IEEE division.
OZ_D float pde_exp(float t, const float*) { return __expf(t); }
OZ_D float pde_rcp(float x) { return __frcp_rn(x); }
OZ_D double pde_rcp(double x) { return fast_rcp(x); }

OZ_D double pde_exp(double t, const double* tab) {
    constexpr double kMagic = 6755399441055744.0;                 // 1.5 * 2^52: adding it rounds to an integer
    constexpr double kLn2Hi = 6.93147180369123816490e-01 / 32;    // ln2 split so that n * kLn2Hi is exact
    constexpr double kLn2Lo = 1.90821492927058770002e-10 / 32;
    const double kd = fma(t, 46.166241308446828384, kMagic);      // 32 log2(e) t
    const int n = __double2loint(kd);
    const double nf = kd - kMagic;
    double r = fma(nf, -kLn2Hi, t);
    r = fma(nf, -kLn2Lo, r);
    double p = fma(r, 1.0 / 720, 1.0 / 120);
    p = fma(p, r, 1.0 / 24);
    p = fma(p, r, 1.0 / 6);
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    const double v = tab[n & (kExpTab - 1)] * p;                  // in [1, 2.1): a NaN argument arrives here as NaN
    // 2^k as two factors, branch-free: overflow, underflow and NaN come out right, and the four exponentials a thread
    // evaluates per stage stay in one basic block (a fast path with an exponent add behind a range test cost a
    // BSSY / BRA / BSYNC triple per call and kept the scheduler from interleaving them)
    const int kc = max(min(n >> 5, 2046), -2044), k1 = kc >> 1;   // both factors stay normal numbers; beyond: inf / 0
    return v * __hiloint2double((k1 + 1023) << 20, 0) * __hiloint2double((kc - k1 + 1023) << 20, 0);
}

// Per-CTA view of the tile.  EDGE = false is the fast path taken by CTAs whose extended tile lies inside the
// domain and does not straddle an end of the control window: no per-cell predicates at all.
template <class M, bool EDGE, int CPT>
struct PdeTile {
    using T = typename M::T;
    bool inside[CPT];
    T uh[CPT];     // 1 where the control acts on the cell, else 0 (EDGE only)
    T uh_all;          // the same for the whole tile (fast path)
    T coef[CPT][M::NCOEF > 0 ? M::NCOEF : 1];   // the model's coefficient fields at this thread's cells
    OZ_D bool in(int m) const { return EDGE ? inside[m] : true; }
    OZ_D T mask(int m) const { return M::WINDOW ? (EDGE ? uh[m] : uh_all) : T(0); }
};

// stages of one explicit step on the CTA's extended tile: F[i][m] for the cells this thread owns.
// `line2` holds two lines used alternately: stage i+1 may overwrite its line while slower warps still read
// stage i's, so one barrier per stencil application is enough.
template <class M, int S, bool WITH_DIAG, bool EDGE, int CPT>
OZ_D void pde_stages(const PdeArgs<M>& a, const PdeTile<M, EDGE, CPT>& tl, typename M::T h, const typename M::T* ulo,
                     const typename M::T* uhi, const typename M::T* yv, typename M::T (*F)[CPT], typename M::T (*Dg)[CPT],
                     typename M::T* line2, const typename M::T* etab, int phase = 0) {
    using T = typename M::T;
    constexpr int P1 = M::P > 0 ? M::P : 1;
    const Tableau<T>& tab = a.tab;
    T du[P1];
#pragma unroll
    for (int q = 0; q < P1; ++q) du[q] = uhi[q] - ulo[q];
#pragma unroll
    for (int i = 0; i < S; ++i) {
        T* line = line2 + ((i + phase) & 1) * (kPdeBlock * CPT + 2);   // phase: stages done before this call (keeps the lines alternating)
        T Yi[CPT];
        T bu_ui[P1];                                       // the controls at the stage abscissa (times beta_U for the built-in model)
#pragma unroll
        for (int q = 0; q < P1; ++q) bu_ui[q] = a.rhs.stage_u(ulo[q], du[q], tab.C[i]);
#pragma unroll
        for (int m = 0; m < CPT; ++m) Yi[m] = yv[m];
#pragma unroll
        for (int j = 0; j < i; ++j) {
            const T haij = h * tab.a(i, j);          // CTA-uniform; the zero entries of a sparse tableau (RK4: 3 of 6)
            if (haij != T(0)) {                      // skip their CPT FMAs with one uniform branch
#pragma unroll
                for (int m = 0; m < CPT; ++m) Yi[m] += haij * F[j][m];
            }
        }
#pragma unroll
        for (int m = 0; m < CPT; ++m) {
            if (EDGE && !tl.inside[m]) Yi[m] = T(0);
            line[1 + threadIdx.x + m * kPdeBlock] = Yi[m];
        }
        __syncthreads();
#pragma unroll
        for (int m = 0; m < CPT; ++m) {
            const int e = threadIdx.x + m * kPdeBlock;
            T dg = T(0);
            T f = a.rhs.template f<WITH_DIAG>(line[e], Yi[m], line[e + 2], tl.mask(m), tl.coef[m], bu_ui, etab, dg);
            if (EDGE && !tl.inside[m]) f = T(0);
            F[i][m] = f;
            if constexpr (WITH_DIAG) Dg[i][m] = dg;
        }
    }
}

template <class M, bool EDGE, int CPT>
OZ_D void pde_tile_setup(const PdeArgs<M>& a, long long base, PdeTile<M, EDGE, CPT>& tl) {
    using T = typename M::T;
    if constexpr (EDGE) {
#pragma unroll
        for (int m = 0; m < CPT; ++m) {
            const long long j = base + threadIdx.x + m * kPdeBlock;
            tl.inside[m] = j >= 0 && j < a.nx;
            if constexpr (M::WINDOW) {
                const long long g = j + a.x_off;
                tl.uh[m] = (g >= a.rhs.h_lo && g < a.rhs.h_hi) ? T(1) : T(0);
            }
        }
    } else if constexpr (M::WINDOW) {
        const long long g = base + a.x_off;
        tl.uh_all = (g >= a.rhs.h_lo && g < a.rhs.h_hi) ? T(1) : T(0);
    }
    if constexpr (M::NCOEF > 0) {
#pragma unroll
        for (int m = 0; m < CPT; ++m) {
            const long long j = base + threadIdx.x + m * kPdeBlock;
            const bool in = j >= 0 && j < a.nx;
#pragma unroll
            for (int q = 0; q < M::NCOEF; ++q) tl.coef[m][q] = in ? __ldg(a.coef + q * a.coef_stride + j + a.x_off) : T(0);
        }
    }
}

// true when the CTA needs the predicated path: it touches an end of the domain or of the control window
template <class M, int CPT>
OZ_D bool pde_is_edge(const PdeArgs<M>& a, long long base) {
    bool cut = false;
    if constexpr (M::WINDOW) {
        const long long g0 = base + a.x_off, g1 = g0 + (kPdeBlock * CPT);   // global range of the extended tile
        const bool cut_lo = g0 < a.rhs.h_lo && g1 > a.rhs.h_lo, cut_hi = g0 < a.rhs.h_hi && g1 > a.rhs.h_hi;
        cut = cut_lo || cut_hi;
    }
    return base < 0 || base + (kPdeBlock * CPT) > a.nx || cut;
}

// Forward launch: a.nsub consecutive steps on the CTA's extended tile.  Every step invalidates s more cells at
// either end of the tile (their neighbours were not part of it), so with a halo of nsub * s the interior is exact
// after the last one; what the fringe computes meanwhile is never stored.  The state stays in registers between the
// fused steps: one global read of the tile and one write of the interior per launch instead of per step, and the
// launch / fill / drain cost of a step-sized kernel (~5 us against ~10 us of work for 10^6 cells) is paid once.
template <class M, int S, bool EDGE, int CPT>
OZ_D void pde_step_body(const PdeArgs<M>& a, long long base, typename M::T* line2, typename M::T* red, const typename M::T* etab) {
    using T = typename M::T;
    const int halo = a.halo, interior = a.tile;
    PdeTile<M, EDGE, CPT> tl;
    pde_tile_setup<M, EDGE, CPT>(a, base, tl);
    T yv[CPT], F[S][CPT];
#pragma unroll
    for (int m = 0; m < CPT; ++m) yv[m] = tl.in(m) ? __ldg(a.y_in + base + threadIdx.x + m * kPdeBlock) : T(0);
    for (int j = 0; j < a.nsub; ++j) {
        const T h = __ldg(a.h + j);
        T ulo[M::P > 0 ? M::P : 1], uhi[M::P > 0 ? M::P : 1];
        ulo[0] = uhi[0] = T(0);
#pragma unroll
        for (int q = 0; q < M::P; ++q) {
            ulo[q] = __ldg(a.u_lo + j * M::P + q);
            uhi[q] = __ldg(a.u_hi + j * M::P + q);
        }
        pde_stages<M, S, false, EDGE, CPT>(a, tl, h, ulo, uhi, yv, F, nullptr, line2, etab, j * S);
        const bool last = j == a.nsub - 1;
        T* dst = a.rows_out ? a.rows_out + j * a.row_stride : (last ? a.y_out : nullptr);
        T acc2 = T(0);
#pragma unroll
        for (int m = 0; m < CPT; ++m) {
            const int e = threadIdx.x + m * kPdeBlock;
            const long long c = base + e;
            T acc = T(0);
#pragma unroll
            for (int i = 0; i < S; ++i) acc += a.tab.B[i] * F[i][m];
            yv[m] += h * acc;                      // F is zero outside the domain: those cells stay zero
            if (e >= halo && e < halo + interior && tl.in(m)) {
                if (dst) dst[c] = yv[m];
                if (c >= a.own_lo && c < a.own_hi) acc2 += a.rhs.q(yv[m], tl.coef[m]);
            }
        }
        if (a.part_sq) {
            const T s = block_sum(acc2, red);
            if (threadIdx.x == 0) a.part_sq[j * a.sq_stride + blockIdx.x] = s;
        }
    }
}

#ifndef OZ_MB_PDE
#define OZ_MB_PDE 3
#endif
// resident CTAs per SM the kernels are compiled for, by cells per thread (fewer cells = fewer registers)
constexpr int pde_mb_fwd(int cpt) { return cpt <= 2 ? 4 : OZ_MB_PDE; }
constexpr int pde_mb_adj(int cpt) { return cpt <= 2 ? 3 : OZ_MB_PDE_ADJ; }

template <class M, int S, int CPT>
__global__ void __launch_bounds__(kPdeBlock, pde_mb_fwd(CPT)) pde_step_kernel(const PdeArgs<M> a) {
    using T = typename M::T;
    __shared__ T line2[2 * (kPdeBlock * CPT + 2)];
    __shared__ T red[32];
    __shared__ T etab[kExpTab];
    pde_exp_table(etab);   // visible after the first barrier inside pde_stages
    const long long base = (long long)blockIdx.x * a.tile - a.halo;
    if (threadIdx.x == 0) {
        line2[0] = line2[(kPdeBlock * CPT + 2)] = T(0);
        line2[(kPdeBlock * CPT) + 1] = line2[(kPdeBlock * CPT + 2) + (kPdeBlock * CPT) + 1] = T(0);
    }
    if (pde_is_edge<M, CPT>(a, base))
        pde_step_body<M, S, true, CPT>(a, base, line2, red, etab);
    else
        pde_step_body<M, S, false, CPT>(a, base, line2, red, etab);
}

// sum of squares of a vector over the owned cells (J_space of the initial condition)
template <class M>
__global__ void __launch_bounds__(kPdeBlock) pde_sumsq_kernel(const PdeArgs<M> a, const typename M::T* __restrict__ y,
                                                              typename M::T* __restrict__ part) {
    using T = typename M::T;
    __shared__ T red[32];
    T acc = T(0);
    for (long long j = a.own_lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; j < a.own_hi;
         j += (long long)gridDim.x * blockDim.x) {
        T cf[M::NCOEF > 0 ? M::NCOEF : 1];
#pragma unroll
        for (int q = 0; q < M::NCOEF; ++q) cf[q] = a.coef[q * a.coef_stride + j + a.x_off];
        acc += a.rhs.q(y[j], cf);
    }
    const T s = block_sum(acc, red);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}

// Reverse launch: a.nsub consecutive reverse steps on the CTA's extended tile, multipliers kept in registers between
// them.  Each step loads the exact y_k of the whole tile, re-derives its stages (s fringe cells per side go stale) and
// runs the transposed chain (s more), so the multipliers of the interior are exact with a halo of 2 s nsub.
template <class M, int S, bool EDGE, int CPT>
OZ_D void pde_adjoint_body(const PdeArgs<M>& a, long long base, typename M::T* line2, typename M::T* red,
                           const typename M::T* etab) {
    using T = typename M::T;
    const Tableau<T>& tab = a.tab;
    const int halo = a.halo, interior = a.tile;
    PdeTile<M, EDGE, CPT> tl;
    pde_tile_setup<M, EDGE, CPT>(a, base, tl);
    T yv[CPT], lam[CPT], F[S][CPT], Yb[S][CPT], Dg[S][CPT];
    bool owned[CPT];
#pragma unroll
    for (int m = 0; m < CPT; ++m) {
        const int e = threadIdx.x + m * kPdeBlock;
        const long long c = base + e;
        lam[m] = tl.in(m) ? __ldg(a.lam_in + c) : T(0);
        owned[m] = e >= halo && e < halo + interior && tl.in(m) && c >= a.own_lo && c < a.own_hi;
    }
    const T fbs = a.field_bar ? __ldg(a.field_bar) * a.fb_scale : T(0);
    for (int j = 0; j < a.nsub; ++j) {
        const T* yrow = a.y_in - j * a.row_stride;
#pragma unroll
        for (int m = 0; m < CPT; ++m) yv[m] = tl.in(m) ? __ldg(yrow + base + threadIdx.x + m * kPdeBlock) : T(0);
        const T h = __ldg(a.h - j);
        constexpr int P1 = M::P > 0 ? M::P : 1;
        T ulo[P1], uhi[P1];
        ulo[0] = uhi[0] = T(0);
#pragma unroll
        for (int q = 0; q < M::P; ++q) {
            ulo[q] = __ldg(a.u_lo - j * M::P + q);
            uhi[q] = __ldg(a.u_hi - j * M::P + q);
        }
        pde_stages<M, S, true, EDGE, CPT>(a, tl, h, ulo, uhi, yv, F, Dg, line2, etab);

        // transposed stage chain:  W_i = b_i lam + sum_{j>i} a_ji Ybar_j,  G_i = h W_i,  Ybar_i = J(Y_i)^T G_i
        // h_bar = sum_i W_i . F_i (no division by h);  u_bar parts from beta_U H . G_i
        T u_lo_acc[P1], u_hi_acc[P1], h_acc = T(0);
#pragma unroll
        for (int q = 0; q < P1; ++q) u_lo_acc[q] = u_hi_acc[q] = T(0);
#pragma unroll
        for (int i = S - 1; i >= 0; --i) {
            T* line = line2 + ((S + (S - 1 - i)) & 1) * (kPdeBlock * CPT + 2);   // keeps alternating after the S forward stages
            T G[CPT];
#pragma unroll
            for (int m = 0; m < CPT; ++m) G[m] = tab.B[i] * lam[m];
#pragma unroll
            for (int q = i + 1; q < S; ++q) {
                const T aqi = tab.a(q, i);
                if (aqi != T(0)) {
#pragma unroll
                    for (int m = 0; m < CPT; ++m) G[m] += aqi * Yb[q][m];
                }
            }
            T usum[P1];
#pragma unroll
            for (int q = 0; q < P1; ++q) usum[q] = T(0);
#pragma unroll
            for (int m = 0; m < CPT; ++m) {
                if (owned[m]) h_acc += G[m] * F[i][m];
                G[m] = (EDGE && !tl.inside[m]) ? T(0) : h * G[m];
                if (owned[m])
#pragma unroll
                    for (int q = 0; q < M::P; ++q) usum[q] += a.rhs.du(tl.mask(m), tl.coef[m], q) * G[m];
                line[1 + threadIdx.x + m * kPdeBlock] = G[m];
            }
            __syncthreads();
#pragma unroll
            for (int m = 0; m < CPT; ++m) {
                const int e = threadIdx.x + m * kPdeBlock;
                const T v = a.rhs.jt(line[e], G[m], line[e + 2], Dg[i][m], tl.coef[m]);
                Yb[i][m] = (EDGE && !tl.inside[m]) ? T(0) : v;
            }
#pragma unroll
            for (int q = 0; q < M::P; ++q) {
                usum[q] *= a.rhs.du_scale();
                u_hi_acc[q] += tab.C[i] * usum[q];
                u_lo_acc[q] += usum[q] - tab.C[i] * usum[q];
            }
        }
        const bool last = j == a.nsub - 1;
        const T* hb = a.hist_bar_k ? a.hist_bar_k - j * a.nx : nullptr;
#pragma unroll
        for (int m = 0; m < CPT; ++m) {
            const int e = threadIdx.x + m * kPdeBlock;
            const long long c = base + e;
            T yb = T(0);
#pragma unroll
            for (int i = 0; i < S; ++i) yb += Yb[i][m];
            T out = lam[m] + yb + fbs * a.rhs.dq(yv[m], tl.coef[m]);
            if (hb && tl.in(m)) out += __ldg(hb + c);
            lam[m] = out;                           // stays zero outside the domain
            if (last && e >= halo && e < halo + interior && tl.in(m)) a.lam_out[c] = out;
        }
        const long long row = -(long long)j * a.adj_stride + blockIdx.x;
        T s = block_sum(h_acc, red);
        if (threadIdx.x == 0) a.part_h[row] = s;
#pragma unroll
        for (int q = 0; q < M::P; ++q) {
            s = block_sum(u_lo_acc[q], red);
            if (threadIdx.x == 0) a.part_ulo[q * a.part_pstride + row] = s;
            s = block_sum(u_hi_acc[q], red);
            if (threadIdx.x == 0) a.part_uhi[q * a.part_pstride + row] = s;
        }
    }
}

template <class M, int S, int CPT>
__global__ void __launch_bounds__(kPdeBlock, pde_mb_adj(CPT)) pde_adjoint_step_kernel(const PdeArgs<M> a) {
    using T = typename M::T;
    __shared__ T line2[2 * (kPdeBlock * CPT + 2)];
    __shared__ T red[32];
    __shared__ T etab[kExpTab];
    pde_exp_table(etab);
    const long long base = (long long)blockIdx.x * a.tile - a.halo;
    if (threadIdx.x == 0) {
        line2[0] = line2[(kPdeBlock * CPT + 2)] = T(0);
        line2[(kPdeBlock * CPT) + 1] = line2[(kPdeBlock * CPT + 2) + (kPdeBlock * CPT) + 1] = T(0);
    }
    if (pde_is_edge<M, CPT>(a, base))
        pde_adjoint_body<M, S, true, CPT>(a, base, line2, red, etab);
    else
        pde_adjoint_body<M, S, false, CPT>(a, base, line2, red, etab);
}

// lam_n = final_bar + hist_bar[n] + field_bar * fb_scale * dq(y_n)      (built-in model: (200/nx) y_n)
template <class M>
__global__ void pde_terminal_kernel(const PdeArgs<M> a, const typename M::T* __restrict__ yn,
                                    const typename M::T* __restrict__ final_bar, const typename M::T* __restrict__ hist_bar_n,
                                    const typename M::T* __restrict__ field_bar, typename M::T* __restrict__ lam) {
    using T = typename M::T;
    const long long nx = a.nx;
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nx) return;
    T cf[M::NCOEF > 0 ? M::NCOEF : 1];
#pragma unroll
    for (int q = 0; q < M::NCOEF; ++q) cf[q] = a.coef[q * a.coef_stride + j + a.x_off];
    T v = (field_bar && yn) ? __ldg(field_bar) * a.fb_scale * a.rhs.dq(yn[j], cf) : T(0);
    if (final_bar) v += final_bar[j];
    if (hist_bar_n) v += hist_bar_n[j];
    lam[j] = v;
}

// field = sum_k [ field_scale sum_cta part_sq[k][cta] + ucost(u_k) ]   (one block, fixed order; built-in: (100/nx), 20 u^2)
template <class M>
__global__ void __launch_bounds__(256) pde_field_kernel(const M rhs, const typename M::T* __restrict__ part_sq, int k0, int nt,
                                                        long long ncta, const typename M::T* __restrict__ u, typename M::T scale,
                                                        typename M::T own_u, typename M::T* __restrict__ field) {
    using T = typename M::T;
    __shared__ T red[32];
    T acc = T(0);
    for (long long q = (long long)k0 * ncta + threadIdx.x; q < (long long)nt * ncta; q += blockDim.x) acc += part_sq[q];
    acc *= scale;
    if (M::P > 0)
        for (int k = k0 + threadIdx.x; k < nt; k += blockDim.x) acc += rhs.ucost(own_u, u + (long long)k * M::P);
    const T s = block_sum(acc, red);
    if (threadIdx.x == 0) field[0] = s;
}

// h_bar[k] = sum_cta part_h[k][cta];  u_bar[k] = sum_cta (part_ulo[k] + part_uhi[k-1]) + own_u * 40 u_k field_bar
template <class M>
__global__ void __launch_bounds__(256) pde_grads_kernel(const M rhs, int n, long long ncta, const typename M::T* __restrict__ part_h,
                                                        const typename M::T* __restrict__ part_ulo,
                                                        const typename M::T* __restrict__ part_uhi,
                                                        const typename M::T* __restrict__ u,
                                                        const typename M::T* __restrict__ field_bar, int k_cost_end,
                                                        long long pstride, typename M::T* __restrict__ h_bar,
                                                        typename M::T* __restrict__ u_bar) {
    using T = typename M::T;
    __shared__ T red[32];
    const int k = blockIdx.x;  // 0..n
    const int q = blockIdx.y;  // control (one row of blocks when the model has none: h_bar only)
    constexpr int P1 = M::P > 0 ? M::P : 1;
    part_ulo += q * pstride;
    part_uhi += q * pstride;
    T ah = T(0), au = T(0);
    for (long long c = threadIdx.x; c < ncta; c += blockDim.x) {
        if (k < n) {
            ah += part_h[(long long)k * ncta + c];
            if (M::P > 0) au += part_ulo[(long long)k * ncta + c];
        }
        if (M::P > 0 && k > 0) au += part_uhi[(long long)(k - 1) * ncta + c];
    }
    T s = block_sum(ah, red);
    if (threadIdx.x == 0 && k < n && h_bar && q == 0) h_bar[k] = s;
    s = block_sum(au, red);
    if (M::P > 0)
        if (threadIdx.x == 0 && u_bar)
            u_bar[(long long)k * P1 + q] = s + ((field_bar && k < k_cost_end) ? rhs.ducost(field_bar[0], u + (long long)k * P1, q) : T(0));
}

template <class M, int S>
static void launch_pde_step(int cpt, unsigned grid, cudaStream_t st, const PdeArgs<M>& a) {
    if (cpt == 2)
        pde_step_kernel<M, S, 2><<<grid, kPdeBlock, 0, st>>>(a);
    else if (cpt == 3)
        pde_step_kernel<M, S, 3><<<grid, kPdeBlock, 0, st>>>(a);
    else
        pde_step_kernel<M, S, 4><<<grid, kPdeBlock, 0, st>>>(a);
}
template <class M, int S>
static void launch_pde_adjoint(int cpt, unsigned grid, cudaStream_t st, const PdeArgs<M>& a) {
    if (cpt == 2)
        pde_adjoint_step_kernel<M, S, 2><<<grid, kPdeBlock, 0, st>>>(a);
    else if (cpt == 3)
        pde_adjoint_step_kernel<M, S, 3><<<grid, kPdeBlock, 0, st>>>(a);
    else
        pde_adjoint_step_kernel<M, S, 4><<<grid, kPdeBlock, 0, st>>>(a);
}

struct PdeLayout {
    long long nx, ncta_f, ncta_a, sq_blocks;
    int tile_f, tile_a;
    int fuse_f, fuse_a;   // steps per forward / reverse launch
    int cpt_f, cpt_a;     // cells per thread of the forward / reverse kernel (tile = 256 * cpt cells including halos)
    size_t off_ping, off_pong, off_sq, off_seg, off_lam0, off_lam1, off_ph, off_pulo, off_puhi, total_fwd, total_adj;
};

template <int S>
static PdeLayout pde_layout(const oz_problem* pb, int P) {
    PdeLayout L;
    L.nx = pb->d_override;
    // Full tiles; the hardware scheduler backfills the ragged last wave better than equalised smaller
    // tiles did (measured: 25.9 ms vs 21.5 ms for 1024 forward steps of 10^6 cells).
    auto balance = [&](int max_tile, int slots, int* tile) {
        long long t = max_tile;
        if (t > L.nx) t = L.nx;
#ifdef OZ_PDE_BALANCE
        const long long full = (L.nx + t - 1) / t, waves = (full + slots - 1) / slots;
        t = (L.nx + waves * slots - 1) / (waves * slots);
        if (t < 1) t = 1;
#else
        (void)slots;
#endif
        *tile = (int)t;
        return (L.nx + t - 1) / t;
    };
    // Steps fused per launch.  A launch costs ~5 us of launch + fill + drain whatever its size, a fused step adds s
    // (forward) / 2 s (reverse) redundant cells per side of every tile.  Big grids (many waves of CTAs) do best with
    // 4 / 2; a grid of a wave or two of CTAs -- the slab of a 2 to 8 GPU spatial shard -- is latency-bound per launch
    // and takes the deeper fusion 16 / 8 (tools/pde_scaling.py, profiles/r2_pde_fusion_sweep.txt: 125k cells 3.79 + 10.34
    // -> 3.26 + 8.31 us per forward / reverse step, 500k cells 8.05 + 20.76 -> 7.34 + 18.76; 10^6 cells are best at 4 / 2).
    // OZ_PDE_FUSE_F / OZ_PDE_FUSE_A (environment) override for experiments.
    const bool small = L.nx <= (long long)OZ_PDE_SMALL_CELLS;
    const bool tiny = L.nx <= (long long)OZ_PDE_TINY_CELLS;
    L.fuse_f = tiny ? 8 : (small ? OZ_PDE_FUSE_SMALL : kPdeFuse);
    L.fuse_a = tiny ? 4 : (small ? OZ_PDE_FUSE_ADJ_SMALL : kPdeFuseAdj);
    if (const char* e = getenv("OZ_PDE_FUSE_F")) L.fuse_f = atoi(e);
    if (const char* e = getenv("OZ_PDE_FUSE_A")) L.fuse_a = atoi(e);
    if (L.fuse_f < 1) L.fuse_f = 1;
    if (L.fuse_a < 1) L.fuse_a = 1;
    // Cells per thread (tile = 256 x cpt cells including halos).  4 is best from 250k cells up, also for the reverse
    // kernel although it spills 152 B there (3 cells per thread do not spill but lose more to halos: 34.8 vs 33.7 us per
    // reverse step at 10^6 cells).  A grid of about one wave of 1024-cell tiles (<= OZ_PDE_TINY_CELLS, the slab of an
    // 8-GPU shard) takes 2 cells per thread AND the shallower fusion 8 / 4: ~280 CTAs of 512 cells, two per SM, so that one
    // computes while the other waits at a stage barrier (125.5k cells: 3.16 + 7.54 -> 2.65 + 6.52 us per forward / reverse
    // step; with 16 / 8 fused steps the halos eat the gain: profiles/r2_pde_cells_per_thread_sweep.txt, r2_pde_tiny_sweep.txt).
    L.cpt_f = tiny ? 2 : kPdeCpt;
    L.cpt_a = tiny ? 2 : kPdeCpt;
    if (const char* e = getenv("OZ_PDE_CPT_F")) L.cpt_f = atoi(e);
    if (const char* e = getenv("OZ_PDE_CPT_A")) L.cpt_a = atoi(e);
    if (L.cpt_f != 2 && L.cpt_f != 3) L.cpt_f = 4;
    if (L.cpt_a != 2 && L.cpt_a != 3) L.cpt_a = 4;
    const int ext_f = kPdeBlock * L.cpt_f, ext_a = kPdeBlock * L.cpt_a;
    while (2 * S * L.fuse_f > ext_f / 2) --L.fuse_f;      // keep at least half of every tile interior
    while (4 * S * L.fuse_a > ext_a / 2) --L.fuse_a;
    L.ncta_f = balance(ext_f - 2 * S * L.fuse_f, 0, &L.tile_f);
    L.ncta_a = balance(ext_a - 4 * S * L.fuse_a, 0, &L.tile_a);
    const size_t nx = (size_t)L.nx, nt = (size_t)pb->n + 1;
    const size_t vec = (nx + 31) / 32 * 32;
    // forward: ping, pong, part_sq[nt][ncta_f]
    L.off_ping = 0;
    L.off_pong = vec;
    L.off_sq = 2 * vec;
    L.total_fwd = L.off_sq + nt * (size_t)L.ncta_f;
    // adjoint: segment states, lambda ping/pong, partials [n][ncta_a] x 3  (+ ping/pong for the replay)
    const size_t ni = pb->hist ? 0 : (size_t)(pb->num_inner > 0 ? pb->num_inner : pb->n);
    L.off_seg = 0;
    L.off_lam0 = ni * vec;
    L.off_lam1 = L.off_lam0 + vec;
    L.off_ph = L.off_lam1 + vec;
    L.off_pulo = L.off_ph + (size_t)pb->n * L.ncta_a;
    L.off_puhi = L.off_pulo + (size_t)P * pb->n * L.ncta_a;
    L.total_adj = L.off_puhi + (size_t)P * pb->n * L.ncta_a;
    return L;
}

template <class M>
static PdeArgs<M> pde_base_args(const oz_problem* pb) {
    using T = typename M::T;
    PdeArgs<M> a;
    memset(&a, 0, sizeof(a));
    a.tab = make_tableau<T>(pb);
    a.nx = pb->d_override;
    // slab mode (grid_nx_global > 0): this rank integrates cells [grid_x_off, grid_x_off + nx) of a larger grid
    // and owns [grid_own_lo, grid_own_hi) of them (local indices) in the reductions
    const bool slab = pb->grid_nx_global > 0;
    a.x_off = slab ? pb->grid_x_off : 0;
    a.nx_global = slab ? pb->grid_nx_global : a.nx;
    a.own_lo = slab ? pb->grid_own_lo : 0;
    a.own_hi = slab ? pb->grid_own_hi : a.nx;
    a.coef = (const T*)pb->th;                       // a traced model's coefficient fields travel as its static parameters
    a.coef_stride = a.nx_global;
    a.rhs.configure(pb, a.nx_global, &a.field_scale, &a.fb_scale);   // model constants; field_scale / fb_scale as above
    return a;
}

template <class M, int S>
static int pde_forward(const oz_problem* pb) {
    using T = typename M::T;
    OZ_REQUIRE(M::NCOEF == 0 || pb->th, "this grid right-hand side needs its coefficient fields (th)");
    OZ_REQUIRE(M::P == 0 || pb->p, "this grid right-hand side needs its dynamic parameters (p)");
    const PdeLayout L = pde_layout<S>(pb, M::P);
    OZ_REQUIRE((size_t)L.total_fwd * sizeof(T) <= pb->ws_bytes && pb->ws, "grid forward workspace too small: need %zu bytes, got %zu",
               (size_t)L.total_fwd * sizeof(T), pb->ws_bytes);
    cudaStream_t st = (cudaStream_t)pb->stream;
    T* ws = (T*)pb->ws;
    const long long nx = L.nx;
    const int n = pb->n, ni = pb->num_inner;
    PdeArgs<M> a = pde_base_args<M>(pb);
    a.tile = L.tile_f;
    const T* h = (const T*)pb->h;
    const T* u = (const T*)pb->p;
    T* hist = (T*)pb->hist;
    T* ckpt = (T*)pb->ckpt;
    T* part_sq = pb->field ? ws + L.off_sq : nullptr;
    const T* cur = (const T*)pb->y0;
    if (hist) OZ_CUDA_CHECK(cudaMemcpyAsync(hist, cur, nx * sizeof(T), cudaMemcpyDeviceToDevice, st));
    if (ckpt) OZ_CUDA_CHECK(cudaMemcpyAsync(ckpt, cur, nx * sizeof(T), cudaMemcpyDeviceToDevice, st));
    if (part_sq) {
        OZ_CUDA_CHECK(cudaMemsetAsync(part_sq, 0, (size_t)L.ncta_f * sizeof(T), st));
        pde_sumsq_kernel<M><<<(unsigned)(L.ncta_f < 1024 ? L.ncta_f : 1024), kPdeBlock, 0, st>>>(a, cur, part_sq);
    }
    a.halo = S * L.fuse_f;
    a.sq_stride = L.ncta_f;
    a.row_stride = nx;
    int launch = 0;
    for (int k = 0; k < n;) {
        // steps k .. k + nsub - 1 in one launch; a group never crosses a checkpoint
        int nsub = n - k < L.fuse_f ? n - k : L.fuse_f;
        if (ckpt && ni > 0 && nsub > ni - k % ni) nsub = ni - k % ni;
        const int kl = k + nsub;                       // index of the state the launch ends on
        const bool at_ckpt = ckpt && ni > 0 && kl % ni == 0;
        T* out;
        if (hist)
            out = hist + (size_t)kl * nx;
        else if (at_ckpt)
            out = ckpt + (size_t)(kl / ni) * nx;
        else if (kl == n && pb->yfinal)
            out = (T*)pb->yfinal;
        else
            out = ws + ((launch & 1) ? L.off_pong : L.off_ping);
        a.nsub = nsub;
        a.y_in = cur;
        a.y_out = hist ? nullptr : out;
        a.rows_out = hist ? hist + (size_t)(k + 1) * nx : nullptr;   // history: every fused step is a row
        a.h = h + k;
        a.u_lo = u + (size_t)k * M::P;
        a.u_hi = u + (size_t)(k + 1) * M::P;
        a.part_sq = part_sq ? part_sq + (size_t)(k + 1) * L.ncta_f : nullptr;
        launch_pde_step<M, S>(L.cpt_f, (unsigned)L.ncta_f, st, a);
        if (hist && at_ckpt)
            OZ_CUDA_CHECK(cudaMemcpyAsync(ckpt + (size_t)(kl / ni) * nx, out, nx * sizeof(T), cudaMemcpyDeviceToDevice, st));
        cur = out;
        k = kl;
        ++launch;
    }
    OZ_CUDA_CHECK(cudaGetLastError());
    if (pb->yfinal && cur != (const T*)pb->yfinal)
        OZ_CUDA_CHECK(cudaMemcpyAsync(pb->yfinal, cur, nx * sizeof(T), cudaMemcpyDeviceToDevice, st));
    if (pb->field) {
        pde_field_kernel<M><<<1, 256, 0, st>>>(a.rhs, part_sq, (pb->grid_flags & OZ_GRID_SKIP_IC_FIELD) ? 1 : 0, n + 1, L.ncta_f, u,
                                               a.field_scale,
                                               (T)((pb->grid_flags & OZ_GRID_NO_CONTROL_COST) ? 0.0 : 1.0), (T*)pb->field);
        OZ_CUDA_CHECK(cudaGetLastError());
    }
    if (pb->status) OZ_CUDA_CHECK(cudaMemsetAsync(pb->status, 0, sizeof(int), st));
    if (pb->iters) OZ_CUDA_CHECK(cudaMemsetAsync(pb->iters, 0, sizeof(int), st));
    return 0;
}

template <class M, int S>
static int pde_adjoint(const oz_problem* pb) {
    using T = typename M::T;
    OZ_REQUIRE(M::NCOEF == 0 || pb->th, "this grid right-hand side needs its coefficient fields (th)");
    const PdeLayout L = pde_layout<S>(pb, M::P);
    OZ_REQUIRE((size_t)L.total_adj * sizeof(T) <= pb->ws_bytes && pb->ws, "grid adjoint workspace too small: need %zu bytes, got %zu",
               (size_t)L.total_adj * sizeof(T), pb->ws_bytes);
    OZ_REQUIRE(pb->hist || pb->ckpt, "adjoint needs hist or ckpt");
    cudaStream_t st = (cudaStream_t)pb->stream;
    T* ws = (T*)pb->ws;
    const long long nx = L.nx;
    const size_t vec = ((size_t)nx + 31) / 32 * 32;
    const int n = pb->n;
    const bool use_seg = pb->hist == nullptr;
    const int ni = use_seg ? (pb->num_inner > 0 ? pb->num_inner : n) : n;
    const int nseg = n / ni;
    PdeArgs<M> a = pde_base_args<M>(pb);
    a.tile = L.tile_a;
    const T* h = (const T*)pb->h;
    const T* u = (const T*)pb->p;
    const T* hist = (const T*)pb->hist;
    const T* ckpt = (const T*)pb->ckpt;
    const T* hist_bar = (const T*)pb->hist_bar;
    T* seg = ws + L.off_seg;
    T* lam[2] = {ws + L.off_lam0, ws + L.off_lam1};
    T* ph = ws + L.off_ph;
    T* pulo = ws + L.off_pulo;
    T* puhi = ws + L.off_puhi;
    a.field_bar = (const T*)pb->field_bar;
    const T* yn = use_seg ? (pb->num_inner > 0 ? ckpt + (size_t)(n / pb->num_inner) * nx : nullptr) : hist + (size_t)n * nx;
    OZ_REQUIRE(yn != nullptr || !pb->field_bar, "a field cotangent needs the final state (hist, or ckpt with num_inner > 0)");
    const unsigned gb = (unsigned)((nx + 255) / 256);
    pde_terminal_kernel<M><<<gb, 256, 0, st>>>(a, yn, (const T*)pb->final_bar, hist_bar ? hist_bar + (size_t)n * nx : nullptr,
                                              (pb->grid_flags & OZ_GRID_SKIP_TERMINAL_FIELD) ? nullptr : a.field_bar, lam[0]);
    int cur = 0;
    for (int o = nseg - 1; o >= 0; --o) {
        const int lo = o * ni;
        if (use_seg) {
            // replay the segment from its checkpoint (time_marching_uniform.py:77-125)
            const T* start = (pb->num_inner > 0) ? ckpt + (size_t)o * nx : (const T*)pb->y0;
            OZ_CUDA_CHECK(cudaMemcpyAsync(seg, start, nx * sizeof(T), cudaMemcpyDeviceToDevice, st));
            for (int q = 0; q + 1 < ni;) {             // rows q+1 .. q+nsub of the segment from row q, fused
                const int nsub = ni - 1 - q < L.fuse_f ? ni - 1 - q : L.fuse_f;
                PdeArgs<M> f = a;
                f.tile = L.tile_f;
                f.halo = S * L.fuse_f;
                f.nsub = nsub;
                f.y_in = seg + (size_t)q * vec;
                f.y_out = nullptr;
                f.rows_out = seg + (size_t)(q + 1) * vec;
                f.row_stride = (long long)vec;
                f.h = h + lo + q;
                f.u_lo = u + (size_t)(lo + q) * M::P;
                f.u_hi = u + (size_t)(lo + q + 1) * M::P;
                f.part_sq = nullptr;
                launch_pde_step<M, S>(L.cpt_f, (unsigned)L.ncta_f, st, f);
                q += nsub;
            }
        }
        for (int k = lo + ni - 1; k >= lo;) {          // steps k .. k - nsub + 1 in one launch, within the segment
            const int nsub = k - lo + 1 < L.fuse_a ? k - lo + 1 : L.fuse_a;
            const int kl = k - nsub + 1;
            a.nsub = nsub;
            a.halo = 2 * S * L.fuse_a;
            a.row_stride = use_seg ? (long long)vec : nx;
            a.adj_stride = L.ncta_a;
            a.y_in = use_seg ? seg + (size_t)(k - lo) * vec : hist + (size_t)k * nx;
            a.lam_in = lam[cur];
            a.lam_out = (kl == 0 && pb->y0_bar) ? (T*)pb->y0_bar : lam[cur ^ 1];
            a.hist_bar_k = hist_bar ? hist_bar + (size_t)k * nx : nullptr;
            a.h = h + k;
            a.u_lo = u + (size_t)k * M::P;
            a.u_hi = u + (size_t)(k + 1) * M::P;
            a.part_pstride = (long long)n * L.ncta_a;
            a.part_h = ph + (size_t)k * L.ncta_a;
            a.part_ulo = pulo + (size_t)k * L.ncta_a;
            a.part_uhi = puhi + (size_t)k * L.ncta_a;
            launch_pde_adjoint<M, S>(L.cpt_a, (unsigned)L.ncta_a, st, a);
            cur ^= 1;
            k = kl - 1;
        }
    }
    OZ_CUDA_CHECK(cudaGetLastError());
    if (pb->h_bar || pb->p_bar) {
        const int k_cost_end = (pb->grid_flags & OZ_GRID_NO_CONTROL_COST) ? 0 : ((pb->grid_flags & OZ_GRID_SKIP_TERMINAL_FIELD) ? n : n + 1);
        pde_grads_kernel<M><<<dim3(n + 1, M::P > 0 ? M::P : 1), 256, 0, st>>>(a.rhs, n, L.ncta_a, ph, pulo, puhi, u, a.field_bar, k_cost_end,
                                                                           (long long)n * L.ncta_a, (T*)pb->h_bar, (T*)pb->p_bar);
        OZ_CUDA_CHECK(cudaGetLastError());
    }
    return 0;
}

#ifndef OZ_WANT_S
#define OZ_WANT_S(s) true
#endif
// M<T> for one dtype; the OZ_ONLY_* macros of a plugin build restrict what is instantiated
template <class M>
static int pde_by_stage(const oz_problem* pb, int mode, size_t* ws_bytes) {
    using T = typename M::T;
#define OZ_PDE_CASE(SS)                                                              \
    if constexpr (OZ_WANT_S(SS))                                                     \
        if (pb->num_stages == SS) {                                                  \
            if (mode == TM_FORWARD) return pde_forward<M, SS>(pb);                   \
            if (mode == TM_ADJOINT) return pde_adjoint<M, SS>(pb);                   \
            if (mode == TM_ADJOINT_WS) {                                             \
                *ws_bytes = pde_layout<SS>(pb, M::P).total_adj * sizeof(T);                \
                return 0;                                                            \
            }                                                                        \
            *ws_bytes = pde_layout<SS>(pb, M::P).total_fwd * sizeof(T);                    \
            return 0;                                                                \
        }
    OZ_PDE_CASE(1)
    OZ_PDE_CASE(2)
    OZ_PDE_CASE(3)
    OZ_PDE_CASE(4)
    OZ_PDE_CASE(7)
#undef OZ_PDE_CASE
    set_error("no grid kernel for %d stages", pb->num_stages);
    return -1;
}

// checks shared by every grid model; mode: TM_FORWARD, TM_ADJOINT, TM_ADJOINT_WS, 100 = forward workspace query
static int pde_check(const oz_problem* pb, int mode, const EvalArgs* ev) {
    OZ_REQUIRE(mode != TM_RHS_EVAL || ev, "internal: missing eval args");
    if (mode == TM_RHS_EVAL || mode == TM_OUT_VJP || mode == TM_JAC_EVAL || mode == TM_JAC_PATTERN) {
        set_error("node evaluation is not available for a grid right-hand side");
        return -1;
    }
    OZ_REQUIRE(pb->is_explicit, "the grid right-hand side supports explicit methods only "
                                "(an implicit step would be a %d x %d Newton system per step)", pb->d_override, pb->d_override);
    OZ_REQUIRE(pb->N == 1, "the grid right-hand side integrates one trajectory (N = 1)");
    OZ_REQUIRE(pb->d_override >= 1, "d_override (number of cells) must be set");
    OZ_REQUIRE(pb->h_shared, "the grid right-hand side needs a shared step vector");
    return 0;
}

}  // namespace oz
