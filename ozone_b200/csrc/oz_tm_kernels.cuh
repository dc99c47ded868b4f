// Fused general-linear-method time-marching kernels (forward + reverse sweep), one trajectory
// per thread, batch innermost in memory so a warp's 32 trajectories read/write 32 consecutive
// scalars.  State, stage values, Newton iterates and the tableau never touch HBM.
//
//   forward  : ozone/integrators/time_marching.py:6-59, :143-260  (+ time_marching_uniform.py for checkpoints)
//   adjoint  : the reverse sweep csdl derives from that loop (call sites lotka_volterra.py:76-79, run_ode.py:66-69)
#pragma once
#include "oz_common.cuh"
#include "oz_linalg.cuh"
#include "oz_sparse_solve.cuh"

namespace oz {

template <class R>
struct Dims {
    static constexpr int D = R::D, P = R::P, SP = R::S, NP = R::NPROF, NF = R::NFIELD;
    static constexpr int P1 = AtLeast1<P>::value, S1 = AtLeast1<SP>::value, NP1 = AtLeast1<NP>::value,
                         NF1 = AtLeast1<NF>::value;
};

// Implicit stage systems up to this many unknowns (after eliminating quadrature states) are solved in
// registers (reg_solve); larger ones go through the shared-memory LU.
// (fp64: a 6 x 6 augmented system is ~170 registers with the pivoting code; 7 x 7 already spills in context)
template <class R, int S, typename T>
constexpr bool use_reg_solve() {
    return S * (R::D - R::DQ) <= (sizeof(T) == 4 ? 8 : 6);
}

// Occupancy targets (threads per block, minimum resident blocks per SM) per kernel flavour.  Implicit
// kernels that solve in registers run 64-thread blocks; the second number caps registers per thread.
#ifndef OZ_ADJ_UNROLL   /* explicit reverse sweep: step-loop unroll factor (lets the compiler rename instead of rotating registers) */
#define OZ_ADJ_UNROLL 2
#endif
#ifndef OZ_SPARSE_SOLVE   /* 1: structure-aware elimination (oz_sparse_solve.cuh); 0: dense register elimination (A/B baseline) */
#define OZ_SPARSE_SOLVE 1
#endif
#ifndef OZ_FWD_SOLVE_EARLY_EXIT   /* forward sweep, by system size M: measured per size on Van der Pol (GL2 M = 2: branch-free 0.072 ->
                                      0.062 ms; GL4 M = 4: branch-free 0.112 -> 0.129 ms; GL6 M = 6: branch-free 0.220 -> 0.195 ms) */
#define OZ_FWD_SOLVE_EARLY_EXIT(M) ((M) == 4)
#endif
#ifndef OZ_ADJ_SOLVE_EARLY_EXIT   /* reverse sweep: pivot tests leave the elimination early (fewer live registers, see oz_sparse_solve.cuh) */
#define OZ_ADJ_SOLVE_EARLY_EXIT true
#endif
#ifndef OZ_IMPL_BLOCK   /* threads per block of the register-solve implicit kernels */
#define OZ_IMPL_BLOCK 64
#endif
#ifndef OZ_MB_IMPL_FWD
#define OZ_MB_IMPL_FWD 4
#endif
#ifndef OZ_MB_IMPL_ADJ
#define OZ_MB_IMPL_ADJ 4
#endif
#ifndef OZ_MB_IMPL_ADJ_SPARSE   /* reverse sweep of a system whose Newton matrix keeps few entries (oz_sparse_solve.cuh): 128 registers */
#define OZ_MB_IMPL_ADJ_SPARSE 8
#endif
#ifndef OZ_SPARSE_LIVE_LIMIT    /* "few": at most this many general entries after fill-in (Van der Pol / Gauss-Legendre 6 has 27, a dense 6 x 6 has 36) */
#define OZ_SPARSE_LIVE_LIMIT 30
#endif
#ifndef OZ_MB_EXPL_ADJ_F64   /* explicit reverse sweep, fp64: 128 registers (5 blocks = 96 registers spills the preloads) */
#define OZ_MB_EXPL_ADJ_F64 4
#endif
#ifndef OZ_MB_EXPL_ADJ_F32
#define OZ_MB_EXPL_ADJ_F32 6
#endif
#ifndef OZ_MB_EXPL_ADJ_LEAN_F64   /* specialised explicit reverse sweep (tm_adjoint_kernel SPEC != 0) */
#define OZ_MB_EXPL_ADJ_LEAN_F64 5
#endif
#ifndef OZ_MB_EXPL_ADJ_LEAN_F32
#define OZ_MB_EXPL_ADJ_LEAN_F32 6
#endif
template <class R, int S, typename T, bool IMPL>
constexpr int lb_threads() {
    return (IMPL && use_reg_solve<R, S, T>()) ? OZ_IMPL_BLOCK : 128;
}
template <class R, int S, typename T, bool IMPL, bool ADJ>
constexpr int lb_blocks() {
    if constexpr (IMPL && use_reg_solve<R, S, T>()) {
        // The reverse sweep factors one system per step and mostly waits on memory (stage cache, history): with a sparse
        // system it fits 128 registers with little spilling and twice the resident warps win (Van der Pol GL6, 65,536
        // trajectories: 0.153 -> 0.112 ms).  The forward sweep (three Newton iterations per step) loses more to the
        // spills than it gains (0.221 -> 0.250 ms), and dense systems spill heavily either way: both stay at 4 blocks.
        if (ADJ) return (OZ_SPARSE_SOLVE && newton_live_entries<R, S>() <= OZ_SPARSE_LIVE_LIMIT) ? OZ_MB_IMPL_ADJ_SPARSE : OZ_MB_IMPL_ADJ;
        return OZ_MB_IMPL_FWD;
    }
    if (!IMPL && ADJ) return sizeof(T) == 8 ? OZ_MB_EXPL_ADJ_F64 : OZ_MB_EXPL_ADJ_F32;
    return 1;
}

// where the register solve leaves unknown (stage i, active component r)
template <class R, int S>
constexpr int sol_idx(int i, int r) {
#if OZ_SPARSE_SOLVE
    return uidx<R, S>(i, r);
#else
    return i * (R::D - R::DQ) + r;
#endif
}

template <bool B>
struct BoolTag {
    static constexpr bool value = B;
};

// Pull a line towards L1 ahead of its use.  The reverse sweep runs at 2-4 warps per scheduler (the
// stage arrays take 120-255 registers), too few to hide a DRAM round trip behind other warps, and every
// step starts with loads that nothing else depends on: their addresses are known one step ahead.
OZ_D void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// row `row` of a [rows][N] (per trajectory) or [rows] (shared) array
template <typename T>
OZ_D T ld_in(const T* base, long long row, long long N, long long i, int shared) {
    return shared ? __ldg(base + row) : __ldg(base + row * N + i);
}

template <class R, typename T>
OZ_D void load_params(const TmArgs<T>& a, int k, long long i, T* p) {
#pragma unroll
    for (int c = 0; c < R::P; ++c) p[c] = ld_in(a.p, (long long)k * R::P + c, a.N, i, a.p_shared);
}

// stage-abscissa interpolation, ozone/collections.py:277-289
template <class R, typename T, int S>
OZ_D void stage_params(const Tableau<T>& tab, const T* plo, const T* phi, T (*ps)[Dims<R>::P1]) {
#pragma unroll
    for (int c = 0; c < R::P; ++c) {
        const T dp = phi[c] - plo[c];         // p[k](1 - c_s) + p[k+1] c_s = p[k] + c_s (p[k+1] - p[k]): one FMA per stage
#pragma unroll
        for (int s = 0; s < S; ++s) ps[s][c] = plo[c] + tab.C[s] * dp;
    }
}

// ---------------------------------------------------------------------------------------------
// Stage solve.  Explicit: time_marching.py:154-188.  Implicit: Newton on
// R(Y) = Y - (A (x) I) h F(Y) - 1 (x) y, initial value 1 (x) y, residual tested before each update
// (time_marching.py:189-243).  On return F holds f at the final stage values Y.
template <class R, typename T, int S, bool IMPL>
OZ_D void solve_stages(const TmArgs<T>& a, const T* y, T h, const T (*ps)[Dims<R>::P1], const T* th,
                       T (*Y)[R::D], T (*F)[R::D], unsigned char* smem, int& status, int& iters) {
    constexpr int D = R::D;
    const Tableau<T>& tab = a.tab;
    if constexpr (!IMPL) {
        // The first stage of an explicit method is y itself (row 0 of A is empty).  Tableaux whose A has only the
        // first subdiagonal (a.chain: RK4 is the one that matters) skip the structurally zero products.
        if (S >= 3 && a.chain) {
#pragma unroll
            for (int i = 0; i < S; ++i) {
                const T ha = i > 0 ? h * tab.a(i, i - 1) : T(0);
#pragma unroll
                for (int c = 0; c < D; ++c) Y[i][c] = i > 0 ? y[c] + ha * F[i - 1][c] : y[c];
                R::f(Y[i], ps[i], th, F[i]);
            }
        } else {
#pragma unroll
            for (int i = 0; i < S; ++i) {
#pragma unroll
                for (int c = 0; c < D; ++c) {
                    if (i == 0) {
                        Y[i][c] = y[c];
                        continue;
                    }
                    T acc = tab.a(i, 0) * F[0][c];
#pragma unroll
                    for (int j = 1; j < i; ++j) acc += tab.a(i, j) * F[j][c];
                    Y[i][c] = y[c] + h * acc;
                }
                R::f(Y[i], ps[i], th, F[i]);
            }
        }
    } else if constexpr (use_reg_solve<R, S, T>()) {
        // Register path.  States are split into DA active and DQ trailing quadrature states (zero Jacobian
        // columns).  The full Newton system is block triangular, [M_aa 0; M_qa I] dY = -R, so
        //   dY_a = -M_aa^-1 R_a ,   dY_q = -R_q + h sum_j a_ij J_j[q,a] dY_a[j]
        // gives exactly the iterates of the full (S D) Newton while factoring only (S DA)^2 entries.
        constexpr int DA = D - R::DQ, MA = S * DA;
        const T tol2 = a.newton_tol * a.newton_tol;
        T hA[S][S];   // h a_ij: used by the residual and by the Newton matrix of every iteration of this step
#pragma unroll
        for (int i = 0; i < S; ++i)
#pragma unroll
            for (int j = 0; j < S; ++j) hA[i][j] = h * tab.a(i, j);
#pragma unroll
        for (int i = 0; i < S; ++i)
#pragma unroll
            for (int c = 0; c < D; ++c) Y[i][c] = y[c];
        for (int it = 0;; ++it) {
#pragma unroll
            for (int i = 0; i < S; ++i) R::f(Y[i], ps[i], th, F[i]);
            T Rs[S][D];
            T rn2 = T(0);
#pragma unroll
            for (int i = 0; i < S; ++i) {
                T part = T(0);   // per-stage partial sums: S short chains instead of one of S * D dependent FMAs
#pragma unroll
                for (int c = 0; c < D; ++c) {
                    T acc = Y[i][c] - y[c];
#pragma unroll
                    for (int j = 0; j < S; ++j) acc -= hA[i][j] * F[j][c];
                    Rs[i][c] = acc;
                    part += acc * acc;
                }
                rn2 += part;
            }
            // one exit test per iteration: ||R||_2 < tol (compared squared), a NaN residual, or out of iterations
            if (!(rn2 >= tol2) || it >= a.newton_max_iter) {
                if (!(rn2 == rn2))
                    status |= OZ_STATUS_NONFINITE;
                else if (!(rn2 < tol2))
                    status |= OZ_STATUS_NOT_CONVERGED;
                break;
            }
            T Ma[MA][MA], b[MA], Jq[S][R::DQ > 0 ? R::DQ : 1][DA];
            // the Newton matrix, entry by entry, through `put` (dense, stage-major: the cold pivoting path, and the
            // dense register solve when OZ_SPARSE_SOLVE is 0)
            auto assemble = [&](auto&& put) {
#pragma unroll
                for (int j = 0; j < S; ++j) {
                    T J[D * D];
                    R::jac(Y[j], ps[j], th, J);
#pragma unroll
                    for (int i = 0; i < S; ++i) {
                        const T ha = hA[i][j];
#pragma unroll
                        for (int r = 0; r < DA; ++r)
#pragma unroll
                            for (int c = 0; c < DA; ++c)
                                put(i * DA + r, j * DA + c, ((i == j && r == c) ? T(1) : T(0)) - ha * J[r * D + c]);
                    }
                }
            };
            int bad;
#if OZ_SPARSE_SOLVE
            // only the entries the Jacobian pattern allows, in the component-major order of oz_sparse_solve.cuh
            static_for<0, S>([&](auto jc) {
                constexpr int j = decltype(jc)::value;
                T J[D * D];
                R::jac(Y[j], ps[j], th, J);
                static_for<0, S>([&](auto ic) {
                    constexpr int i = decltype(ic)::value;
                    static_for<0, DA * DA>([&](auto rcc) {
                        constexpr int r = decltype(rcc)::value / DA, c = decltype(rcc)::value % DA;
                        if constexpr (jac_nz<R>(r, c))
                            Ma[uidx<R, S>(i, r)][uidx<R, S>(j, c)] = ((i == j && r == c) ? T(1) : T(0)) - hA[i][j] * J[r * D + c];
                    });
                });
#pragma unroll
                for (int q = 0; q < R::DQ; ++q)
#pragma unroll
                    for (int c = 0; c < DA; ++c) Jq[j][q][c] = J[(DA + q) * D + c];
            });
            static_for<0, MA>([&](auto kc) {
                constexpr int i = decltype(kc)::value / DA, r = decltype(kc)::value % DA;
                b[uidx<R, S>(i, r)] = -Rs[i][r];
            });
            bad = sparse_solve<R, S, false, T, OZ_FWD_SOLVE_EARLY_EXIT(MA)>(Ma, b);
#else
            assemble([&](int r, int c, T v) { Ma[r][c] = v; });
#pragma unroll
            for (int j = 0; j < S; ++j) {
                T J[D * D];
                R::jac(Y[j], ps[j], th, J);
#pragma unroll
                for (int q = 0; q < R::DQ; ++q)
#pragma unroll
                    for (int c = 0; c < DA; ++c) Jq[j][q][c] = J[(DA + q) * D + c];
            }
#pragma unroll
            for (int i = 0; i < S; ++i)
#pragma unroll
                for (int r = 0; r < DA; ++r) b[i * DA + r] = -Rs[i][r];
            bad = reg_solve<T, MA>(Ma, b);
#endif
            if (bad == 2) {  // a diagonal entry failed the threshold test: solve again with row exchanges
                T ta[MA * MA], tb[MA];
                assemble([&](int r, int c, T v) { ta[r * MA + c] = v; });
#pragma unroll
                for (int i = 0; i < S; ++i)
#pragma unroll
                    for (int r = 0; r < DA; ++r) tb[i * DA + r] = -Rs[i][r];
                bad = finish_with_pivoting<T, MA>(ta, tb, 0) ? 0 : 1;
                static_for<0, MA>([&](auto kc) {
                    constexpr int k = decltype(kc)::value;
                    b[sol_idx<R, S>(k / DA, k % DA)] = tb[k];
                });
            }
            if (bad) {
                status |= OZ_STATUS_SINGULAR;
                break;
            }
            // dYa[i][r]: the Newton update of the active components, read back in stage order
            T dYa[S][DA];
            static_for<0, MA>([&](auto kc) {
                constexpr int i = decltype(kc)::value / DA, r = decltype(kc)::value % DA;
                dYa[i][r] = b[sol_idx<R, S>(i, r)];
            });
            if constexpr (R::DQ > 0) {
                T wq[S][R::DQ];
#pragma unroll
                for (int j = 0; j < S; ++j)
#pragma unroll
                    for (int q = 0; q < R::DQ; ++q) {
                        T acc = T(0);
#pragma unroll
                        for (int c = 0; c < DA; ++c) acc += Jq[j][q][c] * dYa[j][c];
                        wq[j][q] = acc;
                    }
#pragma unroll
                for (int i = 0; i < S; ++i)
#pragma unroll
                    for (int q = 0; q < R::DQ; ++q) {
                        T acc = -Rs[i][DA + q];
#pragma unroll
                        for (int j = 0; j < S; ++j) acc += hA[i][j] * wq[j][q];
                        Y[i][DA + q] += acc;
                    }
            }
#pragma unroll
            for (int i = 0; i < S; ++i)
#pragma unroll
                for (int r = 0; r < DA; ++r) Y[i][r] += dYa[i][r];
            ++iters;
        }
    } else {
        constexpr int M = S * D;
        using Sys = SmemSystem<T, M>;
        Sys sys = Sys::carve(smem, blockDim.x, threadIdx.x);
#pragma unroll
        for (int i = 0; i < S; ++i)
#pragma unroll
            for (int c = 0; c < D; ++c) Y[i][c] = y[c];
        for (int it = 0;; ++it) {
#pragma unroll
            for (int i = 0; i < S; ++i) R::f(Y[i], ps[i], th, F[i]);
            T rn2 = T(0);
#pragma unroll
            for (int i = 0; i < S; ++i)
#pragma unroll
                for (int c = 0; c < D; ++c) {
                    T acc = T(0);
#pragma unroll
                    for (int j = 0; j < S; ++j) acc += tab.a(i, j) * F[j][c];
                    const T r = Y[i][c] - h * acc - y[c];
                    sys.v(i * D + c) = -r;
                    rn2 += r * r;
                }
            if (!(rn2 == rn2)) {  // NaN: give up, flag
                status |= OZ_STATUS_NONFINITE;
                break;
            }
            if (rn2 < a.newton_tol * a.newton_tol) break;   // ||R||_2 < tol without the square root
            if (it >= a.newton_max_iter) {
                status |= OZ_STATUS_NOT_CONVERGED;
                break;
            }
            // Newton matrix: block (i, j) = delta_ij I - h a_ij J_f(Y_j)
#pragma unroll
            for (int j = 0; j < S; ++j) {
                T J[D * D];
                R::jac(Y[j], ps[j], th, J);
#pragma unroll
                for (int i = 0; i < S; ++i) {
                    const T ha = h * tab.a(i, j);
#pragma unroll
                    for (int r = 0; r < D; ++r)
#pragma unroll
                        for (int c = 0; c < D; ++c)
                            sys.m(i * D + r, j * D + c) = ((i == j && r == c) ? T(1) : T(0)) - ha * J[r * D + c];
                }
            }
            if (!sys.factor()) {
                status |= OZ_STATUS_SINGULAR;
                break;
            }
            sys.solve();
#pragma unroll
            for (int i = 0; i < S; ++i)
#pragma unroll
                for (int c = 0; c < D; ++c) Y[i][c] += sys.v(i * D + c);
            ++iters;
        }
    }
}

// y+ = y + h sum_i b_i F_i   (time_marching.py:247-260, V = 1)
template <class R, typename T, int S>
OZ_D void advance(const Tableau<T>& tab, T* y, T h, const T (*F)[R::D]) {
#pragma unroll
    for (int c = 0; c < R::D; ++c) {
        T acc = T(0);
#pragma unroll
        for (int i = 0; i < S; ++i) acc += tab.B[i] * F[i][c];
        y[c] += h * acc;
    }
}

// =============================================================================================
template <class R, typename T, int S, bool IMPL>
__global__ void __launch_bounds__((lb_threads<R, S, T, IMPL>()), (lb_blocks<R, S, T, IMPL, false>()))
    tm_forward_kernel(const TmArgs<T> a) {
    using DM = Dims<R>;
    constexpr int D = DM::D;
    extern __shared__ __align__(16) unsigned char smem[];
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.N) return;
    const long long N = a.N;

    T y[D], th[DM::S1], plo[DM::P1], phi[DM::P1], ps[S][DM::P1], Y[S][D], F[S][D];
    T prof[DM::NP1], fld[DM::NF1], facc[DM::NF1];
#pragma unroll
    for (int c = 0; c < D; ++c) y[c] = __ldg(a.y0 + c * N + i);
#pragma unroll
    for (int c = 0; c < DM::SP; ++c) th[c] = ld_in(a.th, c, N, i, a.th_shared);
    load_params<R>(a, 0, i, plo);
    int status = 0, iters = 0;

    if (a.hist)
#pragma unroll
        for (int c = 0; c < D; ++c) a.hist[c * N + i] = y[c];
    if (a.ckpt)
#pragma unroll
        for (int c = 0; c < D; ++c) a.ckpt[c * N + i] = y[c];
    constexpr bool HAS_OUT = (DM::NP + DM::NF) > 0;
    if constexpr (HAS_OUT) {  // outputs at the initial condition (time_marching.py:104-112)
        R::out(y, plo, th, prof, facc);
        if (a.prof)
#pragma unroll
            for (int c = 0; c < DM::NP; ++c) a.prof[c * N + i] = prof[c];
    }

    // p[k+1] and h[k] are fetched one step ahead: at 2-4 warps per scheduler a load issued at the top of the step
    // that needs it is a fully exposed DRAM round trip (ncu: the first use was the kernel's top stall)
    T pnext[DM::P1], hnext;
    load_params<R>(a, 1, i, pnext);
    hnext = ld_in(a.h, 0, N, i, a.h_shared);
    for (int k = 0; k < a.n; ++k) {
        const T h = hnext;
#pragma unroll
        for (int c = 0; c < DM::P; ++c) phi[c] = pnext[c];
        if (k + 1 < a.n) {
            load_params<R>(a, k + 2, i, pnext);
            hnext = ld_in(a.h, k + 1, N, i, a.h_shared);
        }
        stage_params<R, T, S>(a.tab, plo, phi, ps);
        solve_stages<R, T, S, IMPL>(a, y, h, ps, th, Y, F, smem, status, iters);
        if constexpr (IMPL) {
            if (a.stages)
#pragma unroll
                for (int s = 0; s < S; ++s)
#pragma unroll
                    for (int c = 0; c < D; ++c) a.stages[(((long long)k * S + s) * D + c) * N + i] = Y[s][c];
        }
        advance<R, T, S>(a.tab, y, h, F);
        if (a.hist)
#pragma unroll
            for (int c = 0; c < D; ++c) a.hist[((long long)(k + 1) * D + c) * N + i] = y[c];
        if (a.ckpt && a.num_inner > 0 && (k + 1) % a.num_inner == 0) {
            const long long o = (k + 1) / a.num_inner;
#pragma unroll
            for (int c = 0; c < D; ++c) a.ckpt[(o * D + c) * N + i] = y[c];
        }
        if constexpr (HAS_OUT) {  // raw grid-point parameters (time_marching.py:30-32)
            R::out(y, phi, th, prof, fld);
            if (a.prof)
#pragma unroll
                for (int c = 0; c < DM::NP; ++c) a.prof[((long long)(k + 1) * DM::NP + c) * N + i] = prof[c];
#pragma unroll
            for (int c = 0; c < DM::NF; ++c) facc[c] += fld[c];
        }
#pragma unroll
        for (int c = 0; c < DM::P; ++c) plo[c] = phi[c];
    }
    if (a.yfinal)
#pragma unroll
        for (int c = 0; c < D; ++c) a.yfinal[c * N + i] = y[c];
    if constexpr (DM::NF > 0) {
        if (a.field)
#pragma unroll
            for (int c = 0; c < DM::NF; ++c) a.field[c * N + i] = facc[c];
    }
    if (a.status) a.status[i] = status;
    if (a.iters) a.iters[i] = iters;
}

// =============================================================================================
// Adjoint scratch layout (units of T):
//   [ rows_h * W | rows_p * W | rows_th * W | segment states ni * D * N ]
// where rows_x = number of scalars of a batch-SHARED input x (0 if per trajectory), W = warps in
// the grid.  Each warp writes its shuffle-reduced partial; reduce_rows_kernel sums the W partials
// in a fixed order, so shared-input gradients are deterministic (no fp atomics).
struct AdjWs {
    long long W, rows_h, rows_p, rows_th, seg_elems;
    long long off_h, off_p, off_th, off_seg, total;
};

template <class R>
inline AdjWs tm_adjoint_ws_layout(const oz_problem* pb, int block) {
    AdjWs w;
    const long long nblocks = (pb->N + block - 1) / block;
    w.W = nblocks * (block / 32);
    w.rows_h = (pb->h_shared && pb->h_bar) ? pb->n : 0;
    w.rows_p = (pb->p_shared && pb->p_bar) ? (long long)(pb->n + 1) * R::P : 0;
    w.rows_th = (pb->th_shared && pb->th_bar) ? R::S : 0;
    w.seg_elems = (pb->num_inner > 0 && !pb->hist) ? (long long)pb->num_inner * R::D * pb->N : 0;
    w.off_h = 0;
    w.off_p = w.off_h + w.rows_h * w.W;
    w.off_th = w.off_p + w.rows_p * w.W;
    w.off_seg = w.off_th + w.rows_th * w.W;
    w.off_seg = (w.off_seg + 31) / 32 * 32;
    w.total = w.off_seg + w.seg_elems;
    return w;
}

template <typename T>
struct AdjWsPtr {
    T* h;
    T* p;
    T* th;
    T* seg;
    long long W;
};

template <typename T>
__global__ void __launch_bounds__(256) reduce_rows_kernel(const T* __restrict__ part, long long W, T* __restrict__ out) {
    __shared__ T sm[256];
    const long long row = blockIdx.x;
    T acc = T(0);
    for (long long w = threadIdx.x; w < W; w += 256) acc += part[row * W + w];
    sm[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[row] = sm[0];
}

// One reverse step.  In: y = y_k, lam = dL/dy_{k+1}.  Out: lam = dL/dy_k (without direct
// cotangents of y_k), hb, pb_lo / pb_hi (contributions to p_bar[k] / p_bar[k+1]), tb +=.
//   G_i = h b_i lam + h sum_j a_ji Ybar_j,  Ybar_j = J_y(Y_j)^T G_j
//   explicit: back-substitution i = S-1..0;  implicit: M^T Ybar = blk(J^T)(h b (x) lam), M the Newton matrix
template <class R, typename T, int S, bool IMPL>
OZ_D void adjoint_step(const TmArgs<T>& a, const T* y, T h, const T (*ps)[Dims<R>::P1], const T* th, T* lam,
                       T& hb, T* pb_lo, T* pb_hi, T* tb, unsigned char* smem, int& status, int& iters,
                       const T* cached_stages, long long N, long long i) {
    using DM = Dims<R>;
    constexpr int D = DM::D;
    const Tableau<T>& tab = a.tab;
    T Y[S][D], F[S][D], Yb[S][D];
    if (IMPL && cached_stages) {  // converged stages of this step, stored by the forward sweep
#pragma unroll
        for (int s = 0; s < S; ++s) {
#pragma unroll
            for (int c = 0; c < D; ++c) Y[s][c] = __ldg(cached_stages + ((long long)s * D + c) * N + i);
            R::f(Y[s], ps[s], th, F[s]);
        }
    } else {
        solve_stages<R, T, S, IMPL>(a, y, h, ps, th, Y, F, smem, status, iters);
    }

    if constexpr (IMPL && use_reg_solve<R, S, T>()) {
        // M^T = [M_aa^T M_qa^T; 0 I] and the quadrature columns of J are zero, so Ybar_q = 0 and
        // M_aa^T Ybar_a = blk(J^T)(h b (x) lam) restricted to the active components.
        constexpr int DA = D - R::DQ, MA = S * DA;
        T Mt[MA][MA], b[MA];
        auto assemble = [&](auto&& put, auto&& put_rhs) {  // transposed Newton matrix and right-hand side, entry by entry
#pragma unroll
            for (int j = 0; j < S; ++j) {
                T J[D * D];
                R::jac(Y[j], ps[j], th, J);
#pragma unroll
                for (int i = 0; i < S; ++i) {
                    const T ha = h * tab.a(i, j);
#pragma unroll
                    for (int r = 0; r < DA; ++r)
#pragma unroll
                        for (int c = 0; c < DA; ++c)
                            put(j * DA + c, i * DA + r, ((i == j && r == c) ? T(1) : T(0)) - ha * J[r * D + c]);
                }
#pragma unroll
                for (int c = 0; c < DA; ++c) {
                    T acc = T(0);
#pragma unroll
                    for (int r = 0; r < D; ++r) acc += J[r * D + c] * lam[r];
                    put_rhs(j * DA + c, h * tab.B[j] * acc);
                }
            }
        };
        int bad;
#if OZ_SPARSE_SOLVE
        static_for<0, S>([&](auto jc) {
            constexpr int j = decltype(jc)::value;
            T J[D * D];
            R::jac(Y[j], ps[j], th, J);
            static_for<0, S>([&](auto ic) {
                constexpr int i = decltype(ic)::value;
                const T ha = h * tab.a(i, j);
                static_for<0, DA * DA>([&](auto rcc) {
                    constexpr int r = decltype(rcc)::value / DA, c = decltype(rcc)::value % DA;
                    if constexpr (jac_nz<R>(r, c))
                        Mt[uidx<R, S>(j, c)][uidx<R, S>(i, r)] = ((i == j && r == c) ? T(1) : T(0)) - ha * J[r * D + c];
                });
            });
            static_for<0, DA>([&](auto cc) {
                constexpr int c = decltype(cc)::value;
                T acc = -T(0);
                static_for<0, D>([&](auto rr) {
                    constexpr int r = decltype(rr)::value;
                    if constexpr (jac_nz<R>(r, c)) acc += J[r * D + c] * lam[r];
                });
                b[uidx<R, S>(j, c)] = h * tab.B[j] * acc;
            });
        });
        bad = sparse_solve<R, S, true, T, OZ_ADJ_SOLVE_EARLY_EXIT>(Mt, b);
#else
        assemble([&](int r, int c, T v) { Mt[r][c] = v; }, [&](int r, T v) { b[r] = v; });
        bad = reg_solve<T, MA>(Mt, b);
#endif
        if (bad == 2) {  // a diagonal entry failed the threshold test: solve again with row exchanges
            T ta[MA * MA], tb[MA];
            assemble([&](int r, int c, T v) { ta[r * MA + c] = v; }, [&](int r, T v) { tb[r] = v; });
            bad = finish_with_pivoting<T, MA>(ta, tb, 0) ? 0 : 1;
            static_for<0, MA>([&](auto kc) {
                constexpr int k = decltype(kc)::value;
                b[sol_idx<R, S>(k / DA, k % DA)] = tb[k];
            });
        }
        if (bad) status |= OZ_STATUS_SINGULAR;
        static_for<0, S>([&](auto ic) {
            constexpr int i = decltype(ic)::value;
            static_for<0, DA>([&](auto cc) {
                constexpr int c = decltype(cc)::value;
                Yb[i][c] = b[sol_idx<R, S>(i, c)];
            });
#pragma unroll
            for (int c = DA; c < D; ++c) Yb[i][c] = T(0);
        });
    } else if constexpr (IMPL) {
        constexpr int M = S * D;
        using Sys = SmemSystem<T, M>;
        Sys sys = Sys::carve(smem, blockDim.x, threadIdx.x);
#pragma unroll
        for (int j = 0; j < S; ++j) {
            T J[D * D];
            R::jac(Y[j], ps[j], th, J);
#pragma unroll
            for (int i = 0; i < S; ++i) {
                const T ha = h * tab.a(i, j);
#pragma unroll
                for (int r = 0; r < D; ++r)
#pragma unroll
                    for (int c = 0; c < D; ++c)
                        sys.m(i * D + r, j * D + c) = ((i == j && r == c) ? T(1) : T(0)) - ha * J[r * D + c];
            }
            // right-hand side block j:  J_j^T (h b_j lam)
#pragma unroll
            for (int c = 0; c < D; ++c) {
                T acc = T(0);
#pragma unroll
                for (int r = 0; r < D; ++r) acc += J[r * D + c] * lam[r];
                sys.v(j * D + c) = h * tab.B[j] * acc;
            }
        }
        if (!sys.factor()) status |= OZ_STATUS_SINGULAR;
        sys.solve_transposed();
#pragma unroll
        for (int i = 0; i < S; ++i)
#pragma unroll
            for (int c = 0; c < D; ++c) Yb[i][c] = sys.v(i * D + c);
    }

    hb = -T(0);
#pragma unroll
    for (int c = 0; c < DM::P; ++c) {
        pb_lo[c] = -T(0);
        pb_hi[c] = -T(0);
    }
    auto stage_sweep = [&](auto chain_tag) {
        [[maybe_unused]] constexpr bool CHAIN = decltype(chain_tag)::value;
    #pragma unroll
        for (int i = S - 1; i >= 0; --i) {
            T G[D], pbi[DM::P1];
    #pragma unroll
            for (int c = 0; c < DM::P1; ++c) pbi[c] = -T(0);  // -0 is the exact additive identity: the first += folds away
    #pragma unroll
            for (int c = 0; c < D; ++c) {
                T acc = tab.B[i] * lam[c];
                if constexpr (IMPL) {
    #pragma unroll
                    for (int j = 0; j < S; ++j) acc += tab.a(j, i) * Yb[j][c];
                } else if constexpr (CHAIN) {
                    if (i + 1 < S) acc += tab.a(i + 1, i) * Yb[i + 1][c];
                } else {
    #pragma unroll
                    for (int j = i + 1; j < S; ++j) acc += tab.a(j, i) * Yb[j][c];
                }
                hb += acc * F[i][c];  // h_bar = sum_i (G_i / h) . F_i without dividing by h (h = 0 is legal)
                G[c] = h * acc;
            }
            if constexpr (IMPL) {
                T dummy[D];
    #pragma unroll
                for (int c = 0; c < D; ++c) dummy[c] = T(0);
                R::vjp(Y[i], ps[i], th, G, dummy, pbi, tb);
            } else {
    #pragma unroll
                for (int c = 0; c < D; ++c) Yb[i][c] = -T(0);
                R::vjp(Y[i], ps[i], th, G, Yb[i], pbi, tb);
            }
    #pragma unroll
            for (int c = 0; c < DM::P; ++c) {
                pb_lo[c] += pbi[c];               // running total; the (1 - c_i) part is total - c_i part, taken below
                pb_hi[c] += tab.C[i] * pbi[c];
            }
        }
    };
    if (!IMPL && S >= 3 && a.chain)
        stage_sweep(BoolTag<true>{});
    else
        stage_sweep(BoolTag<false>{});
#pragma unroll
    for (int c = 0; c < DM::P; ++c) pb_lo[c] -= pb_hi[c];
    // (h_bar: sum_i G_i . F_i = h [sum_i b_i F_i . lam + sum_j Ybar_j . sum_i a_ji F_i], accumulated above as
    //  sum_i (G_i / h) . F_i from the not-yet-scaled cotangent)
#pragma unroll
    for (int c = 0; c < D; ++c) {
        T acc = T(0);
#pragma unroll
        for (int i = 0; i < S; ++i) acc += Yb[i][c];
        lam[c] += acc;
    }
}

// SPEC selects a compile-time specialisation of the reverse sweep for the commonest request of an optimisation loop --
// a cotangent on the final state only, gradients w.r.t. the shared step vector and the parameters, full history
// stored, subdiagonal tableau (RK4):
//   0  generic: every option is read from the argument block at run time
//   1  final-state cotangent only, h shared + h_bar, p shared + p_bar
//   2  final-state cotangent only, h shared + h_bar, p per trajectory + p_bar
// The specialised kernels run the same code; the options are pinned to constants below, so the compiler removes
// the per-step null-pointer branches, the second copy of the stage sweep and the output-cotangent arithmetic
// (ncu, round 1: 12 uniform branches and ~22 register moves per step on a sweep that is issue-bound).
template <class R, typename T, int S, bool IMPL, int SPEC>
OZ_D void tm_adjoint_body(const TmArgs<T>& a, const AdjWsPtr<T>& ws);

// resident blocks the reverse kernel is compiled for: the specialised explicit sweeps carry far less state than the
// generic one (80-94 registers instead of 128 in fp64), so more of them fit
template <class R, typename T, int S, bool IMPL, int SPEC>
constexpr int lb_blocks_adj() {
    if (!IMPL && SPEC != 0) return sizeof(T) == 8 ? OZ_MB_EXPL_ADJ_LEAN_F64 : OZ_MB_EXPL_ADJ_LEAN_F32;
    return lb_blocks<R, S, T, IMPL, true>();
}

template <class R, typename T, int S, bool IMPL, int SPEC = 0>
__global__ void __launch_bounds__((lb_threads<R, S, T, IMPL>()), (lb_blocks_adj<R, T, S, IMPL, SPEC>()))
    tm_adjoint_kernel(const TmArgs<T> a_in, const AdjWsPtr<T> ws) {
    if constexpr (SPEC == 0) {
        tm_adjoint_body<R, T, S, IMPL, 0>(a_in, ws);      // straight from the parameter block (constant bank)
    } else {
        TmArgs<T> a = a_in;
        a.hist_bar = nullptr;
        a.prof_bar = nullptr;
        a.field_bar = nullptr;
        a.num_inner = 0;
        a.chain = 1;
        a.h_shared = 1;
        a.p_shared = SPEC == 1 ? 1 : 0;
        __builtin_assume(a.h_bar != nullptr);
        __builtin_assume(a.p_bar != nullptr);
        __builtin_assume(a.hist != nullptr);
        __builtin_assume(a.final_bar != nullptr);
        tm_adjoint_body<R, T, S, IMPL, SPEC>(a, ws);
    }
}

template <class R, typename T, int S, bool IMPL, int SPEC>
OZ_D void tm_adjoint_body(const TmArgs<T>& a, const AdjWsPtr<T>& ws) {
    using DM = Dims<R>;
    constexpr int D = DM::D;
    constexpr bool HAS_OUT = (DM::NP + DM::NF) > 0;
    extern __shared__ __align__(16) unsigned char smem[];
    const long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = gi < a.N;
    const long long i = valid ? gi : a.N - 1;  // idle lanes shadow the last trajectory and never store
    const long long N = a.N;
    const long long warp = gi >> 5;
    const int lane = threadIdx.x & 31;
    const int n = a.n;

    T th[DM::S1], tb[DM::S1], lam[D], y[D], plo[DM::P1], phi[DM::P1], ps[S][DM::P1];
    T pb_next[DM::P1], pb_lo[DM::P1], pb_hi[DM::P1], vf[DM::NF1], vp[DM::NP1];
    int status = 0, iters = 0;
#pragma unroll
    for (int c = 0; c < DM::SP; ++c) th[c] = ld_in(a.th, c, N, i, a.th_shared);
#pragma unroll
    for (int c = 0; c < DM::S1; ++c) tb[c] = T(0);
#pragma unroll
    for (int c = 0; c < DM::P1; ++c) pb_next[c] = T(0);
#pragma unroll
    for (int c = 0; c < DM::NF; ++c) vf[c] = a.field_bar ? __ldg(a.field_bar + c * N + i) : T(0);

    const bool use_seg = (a.num_inner > 0) && (a.hist == nullptr);
    // Row offsets are carried from step to step (one 64-bit subtraction each) instead of being rebuilt from k:
    // the reverse sweep of a cheap right-hand side is otherwise issue-bound on address arithmetic.
    const long long DN = (long long)D * N, PN = (long long)DM::P * N, NPN = (long long)DM::NP * N;
    const long long PW = (long long)DM::P * ws.W;
    auto emit_p = [&](int k, long long wrow) {  // p_bar[k] is complete; wrow = k * P * W + warp
        if (!a.p_bar) return;
#pragma unroll
        for (int c = 0; c < DM::P; ++c) {
            if (a.p_shared) {
                const T s = warp_sum(valid ? pb_next[c] : T(0));
                if (lane == 0) ws.p[wrow + c * ws.W] = s;
            } else if (valid) {
                a.p_bar[((long long)k * DM::P + c) * N + i] = pb_next[c];
            }
        }
    };
    // direct cotangents of y_k: history row (hb, already loaded), profile/field outputs evaluated at (y_k, p[k])
    auto load_prof_bar = [&](long long oprof, T* v) {
#pragma unroll
        for (int c = 0; c < DM::NP; ++c) v[c] = a.prof_bar ? __ldg(a.prof_bar + oprof + c * N) : T(0);
    };
    auto direct = [&](const T* hb, const T* yk, const T* pk) {  // vp holds prof_bar[k]
        if (a.hist_bar)
#pragma unroll
            for (int c = 0; c < D; ++c) lam[c] += hb[c];
        if constexpr (HAS_OUT && SPEC == 0) R::out_vjp(yk, pk, th, vp, vf, lam, pb_next, tb);
    };
    auto load_hist_bar = [&](long long oy, T* hb) {
        if (a.hist_bar)
#pragma unroll
            for (int c = 0; c < D; ++c) hb[c] = __ldg(a.hist_bar + oy + c * N);
    };
    auto params_at = [&](int k, long long op, T* p) {  // p[k]: row k of [nt][P] (shared) or [nt][P][N]
#pragma unroll
        for (int c = 0; c < DM::P; ++c) p[c] = a.p_shared ? __ldg(a.p + k * DM::P + c) : __ldg(a.p + op + c * N);
    };

    // Sums over the batch (cotangents of shared h and shared p).  Explicit sweeps, whose steps are a few hundred
    // instructions, do not reduce with shuffles every step (5 shuffle pairs + 5 adds per quantity): each warp parks
    // its lanes' values in a shared-memory tile, one row per (step, quantity), and once the tile holds 32 rows lane r
    // adds up row r and writes that warp partial.  Same partial-sum layout as the shuffle path, fixed order, no atomics.
    constexpr bool STAGED = !IMPL && DM::P + 1 <= 32;   // a step never stages more than 1 + P rows
    constexpr int kRow = 33;  // row pitch: lane r reading row r at a fixed column touches distinct banks
    __shared__ T red_tile[STAGED ? (lb_threads<R, S, T, IMPL>() / 32) * 32 * kRow : 1];
    T* tile = red_tile + (STAGED ? (threadIdx.x >> 5) * 32 * kRow : 0);
    const bool h_red = a.h_bar && a.h_shared, p_red = a.p_bar && a.p_shared && DM::P > 0;
    const int nq = (h_red ? 1 : 0) + (p_red ? DM::P : 0);   // rows per step
    const int kc = nq ? 32 / nq : 0;                        // steps per tile
    int fill = 0, kstart = 0;
    auto flush = [&]() {  // rows 0 .. fill * nq - 1 hold steps kstart, kstart - 1, ...
        __syncwarp();
        if (lane < fill * nq) {
            const T* row = tile + lane * kRow;
            T s0 = row[0], s1 = row[1];
#pragma unroll
            for (int j = 2; j < 32; j += 2) {
                s0 += row[j];
                s1 += row[j + 1];
            }
            const int kk = lane / nq, q = lane - kk * nq, k = kstart - kk;
            if (h_red && q == 0)
                ws.h[(long long)k * ws.W + warp] = s0 + s1;
            else
                ws.p[((long long)(k + 1) * DM::P + (q - (h_red ? 1 : 0))) * ws.W + warp] = s0 + s1;
        }
        __syncwarp();
        fill = 0;
    };

    // ---- terminal condition: lam = dL/dy_n
    T hbv[D];
#pragma unroll
    for (int c = 0; c < D; ++c) lam[c] = a.final_bar ? __ldg(a.final_bar + c * N + i) : T(0);
    params_at(n, (long long)n * PN + i, phi);
    if constexpr (HAS_OUT) {
        const T* src = use_seg ? a.ckpt + (long long)(n / a.num_inner) * DN : a.hist + (long long)n * DN;
#pragma unroll
        for (int c = 0; c < D; ++c) y[c] = __ldg(src + c * N + i);
    }
    load_hist_bar((long long)n * DN + i, hbv);
    load_prof_bar((long long)n * NPN + i, vp);
    direct(hbv, y, phi);

    const int ni = use_seg ? a.num_inner : n;
    const int nseg = use_seg ? n / a.num_inner : 1;
    // offsets of row k = n - 1, the first step of the sweep; every iteration moves them one row down
    long long oy = (long long)(n - 1) * DN + i;      // hist / hist_bar row
    long long op = (long long)(n - 1) * PN + i;      // p row
    long long oprof = (long long)(n - 1) * NPN + i;  // prof_bar row
    long long oh = (long long)(n - 1) * N + i;       // h row (per trajectory)
    long long wrow_p = (long long)n * PW + warp;     // ws.p row of p_bar[k + 1]
    long long wrow_h = (long long)(n - 1) * ws.W + warp;
    // Explicit sweeps load every per-step input one iteration before it is used (registers to spare, and a step is
    // too short for five warps per scheduler to cover an L2 round trip): y_k, h_k, p_k here, the cotangent rows
    // of step k at the top of iteration k for use at its end.
    constexpr bool PRELOAD = !IMPL;
    T hnx = T(0), pnx[DM::P1];
    if constexpr (PRELOAD) {
        if (n > 0) {
            hnx = a.h_shared ? __ldg(a.h + (n - 1)) : __ldg(a.h + oh);
            params_at(n - 1, op, pnx);
        }
    }
    for (int o = nseg - 1; o >= 0; --o) {
        const int lo = o * ni;
        if (use_seg) {
            // recompute the inner states of this segment from its checkpoint
            // (time_marching_uniform.py:77-125: the inner loop is not stacked)
            T Y[S][D], F[S][D], qlo[DM::P1], qhi[DM::P1];
#pragma unroll
            for (int c = 0; c < D; ++c) y[c] = __ldg(a.ckpt + ((long long)o * D + c) * N + i);
            load_params<R>(a, lo, i, qlo);
            for (int q = 0; q < ni; ++q) {
                if (valid)
#pragma unroll
                    for (int c = 0; c < D; ++c) ws.seg[((long long)q * D + c) * N + i] = y[c];
                if (q == ni - 1) break;
                const T h = ld_in(a.h, lo + q, N, i, a.h_shared);
                load_params<R>(a, lo + q + 1, i, qhi);
                stage_params<R, T, S>(a.tab, qlo, qhi, ps);
                int st2 = 0, it2 = 0;
                solve_stages<R, T, S, IMPL>(a, y, h, ps, th, Y, F, smem, st2, it2);
                advance<R, T, S>(a.tab, y, h, F);
#pragma unroll
                for (int c = 0; c < DM::P; ++c) qlo[c] = qhi[c];
            }
        }
        // y_k is fetched one iteration ahead so that its load latency overlaps the previous step's arithmetic:
        // into registers for explicit methods, as an L1 prefetch for implicit ones (no registers to spare there)
        T ynext[D];
        constexpr int kStepUnroll = IMPL ? 1 : OZ_ADJ_UNROLL;
        // row of y_k: the stored history, or this segment's recomputed rows (row k - lo of ws.seg)
        const T* yrow = (use_seg ? ws.seg + (long long)(ni - 1) * DN : a.hist + (long long)(lo + ni - 1) * DN) + i;
        if constexpr (PRELOAD) {
#pragma unroll
            for (int c = 0; c < D; ++c) ynext[c] = use_seg ? yrow[c * N] : __ldg(yrow + c * N);
        }
#pragma unroll kStepUnroll
        for (int k = lo + ni - 1; k >= lo; --k) {
            if constexpr (IMPL) {
                // implicit steps are long and register-starved: rebuild the offsets here rather than keep seven
                // 64-bit values alive across the Newton solve and the elimination
                yrow = (use_seg ? ws.seg + (long long)(k - lo) * DN : a.hist + (long long)k * DN) + i;
                oy = (long long)k * DN + i;
                op = (long long)k * PN + i;
                oprof = (long long)k * NPN + i;
                oh = (long long)k * N + i;
                wrow_p = (long long)(k + 1) * PW + warp;
                wrow_h = (long long)k * ws.W + warp;
            }
            if constexpr (PRELOAD) {
#pragma unroll
                for (int c = 0; c < D; ++c) y[c] = ynext[c];
                load_hist_bar(oy, hbv);  // both consumed at the end of this iteration
                if constexpr (DM::NP > 0) load_prof_bar(oprof, vp);
            } else {
#pragma unroll
                for (int c = 0; c < D; ++c) y[c] = use_seg ? yrow[c * N] : __ldg(yrow + c * N);
            }
            if (k > lo) {  // next iteration's inputs (step k-1)
                const T* nrow = yrow - DN;
                if constexpr (PRELOAD) {
#pragma unroll
                    for (int c = 0; c < D; ++c) ynext[c] = use_seg ? nrow[c * N] : __ldg(nrow + c * N);
                } else {
#pragma unroll
                    for (int c = 0; c < D; ++c) prefetch_l1(nrow + c * N);
                    if (a.hist_bar)
#pragma unroll
                        for (int c = 0; c < D; ++c) prefetch_l1(a.hist_bar + (oy - DN) + c * N);
                }
                if constexpr (!PRELOAD) {
                    if (!a.p_shared)
#pragma unroll
                        for (int c = 0; c < DM::P; ++c) prefetch_l1(a.p + (op - PN) + c * N);
                    if (!a.h_shared) prefetch_l1(a.h + (oh - N));
                    if constexpr (DM::NP > 0) {
                        if (a.prof_bar)
#pragma unroll
                            for (int c = 0; c < DM::NP; ++c) prefetch_l1(a.prof_bar + (oprof - NPN) + c * N);
                    }
                }
                if constexpr (IMPL) {
                    if (a.stages && !use_seg)
#pragma unroll
                        for (int c = 0; c < S * D; ++c) prefetch_l1(a.stages + ((long long)(k - 1) * S * D + c) * N + i);
                }
            }
            T h;
            if constexpr (PRELOAD) {
                h = hnx;
#pragma unroll
                for (int c = 0; c < DM::P; ++c) plo[c] = pnx[c];
                if (k > 0) {  // rows of step k - 1 (also across a segment boundary: h and p are global rows)
                    hnx = a.h_shared ? __ldg(a.h + (k - 1)) : __ldg(a.h + (oh - N));
                    params_at(k - 1, op - PN, pnx);
                }
            } else {
                h = a.h_shared ? __ldg(a.h + k) : __ldg(a.h + oh);
                params_at(k, op, plo);
            }
            stage_params<R, T, S>(a.tab, plo, phi, ps);
            T hb;
            adjoint_step<R, T, S, IMPL>(a, y, h, ps, th, lam, hb, pb_lo, pb_hi, tb, smem, status, iters,
                                        (IMPL && a.stages && !use_seg) ? a.stages + (long long)k * S * D * N : nullptr, N, i);
#pragma unroll
            for (int c = 0; c < DM::P; ++c) pb_next[c] += pb_hi[c];
            if constexpr (STAGED) {
                // shared-input cotangents of this step go to the warp's tile, one row per quantity; per-trajectory
                // ones are stored directly
                if (fill == 0) kstart = k;
                T* row = tile + fill * nq * kRow + lane;
                if (h_red) {
                    *row = valid ? hb : T(0);
                    row += kRow;
                } else if (a.h_bar && valid) {
                    a.h_bar[oh] = hb;
                }
                if (p_red) {
#pragma unroll
                    for (int c = 0; c < DM::P; ++c) row[c * kRow] = valid ? pb_next[c] : T(0);
                } else {
                    emit_p(k + 1, wrow_p);
                }
                if (++fill == kc) flush();
            } else {
                emit_p(k + 1, wrow_p);
                if (a.h_bar) {
                    if (a.h_shared) {
                        const T s = warp_sum(valid ? hb : T(0));
                        if (lane == 0) ws.h[wrow_h] = s;
                    } else if (valid) {
                        a.h_bar[oh] = hb;
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < DM::P; ++c) pb_next[c] = pb_lo[c];
            if constexpr (!PRELOAD) {
                load_hist_bar(oy, hbv);
                if constexpr (DM::NP > 0) load_prof_bar(oprof, vp);
            }
            direct(hbv, y, plo);
#pragma unroll
            for (int c = 0; c < DM::P; ++c) phi[c] = plo[c];
            if constexpr (!IMPL) {
                yrow -= DN;
                oy -= DN;
                op -= PN;
                oprof -= NPN;
                oh -= N;
                wrow_p -= PW;
                wrow_h -= ws.W;
            }
        }
    }
    if constexpr (STAGED) {
        if (fill > 0) flush();
    }
    emit_p(0, warp);
    if (a.y0_bar && valid)
#pragma unroll
        for (int c = 0; c < D; ++c) a.y0_bar[c * N + i] = lam[c];
    if (a.th_bar) {
#pragma unroll
        for (int c = 0; c < DM::SP; ++c) {
            if (a.th_shared) {
                const T s = warp_sum(valid ? tb[c] : T(0));
                if (lane == 0) ws.th[(long long)c * ws.W + warp] = s;
            } else if (valid) {
                a.th_bar[c * N + i] = tb[c];
            }
        }
    }
    if (a.status && valid && status) atomicOr(a.status + i, status);
}

// =============================================================================================
// FuncWrapper.evaluate (ozone/func_wrapper.py:90-107) for `nodes` points.
template <class R, typename T>
__global__ void __launch_bounds__(128) rhs_eval_kernel(long long nodes, int th_shared, const T* __restrict__ y,
                                                       const T* __restrict__ p, const T* __restrict__ th,
                                                       T* __restrict__ dy, T* __restrict__ prof, T* __restrict__ field) {
    using DM = Dims<R>;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nodes) return;
    T yy[DM::D], pp[DM::P1], tt[DM::S1], d[DM::D], pr[DM::NP1], fl[DM::NF1];
#pragma unroll
    for (int c = 0; c < DM::D; ++c) yy[c] = y[c * nodes + i];
#pragma unroll
    for (int c = 0; c < DM::P; ++c) pp[c] = p[c * nodes + i];
#pragma unroll
    for (int c = 0; c < DM::SP; ++c) tt[c] = th_shared ? th[c] : th[c * nodes + i];
    R::f(yy, pp, tt, d);
    if (dy)
#pragma unroll
        for (int c = 0; c < DM::D; ++c) dy[c * nodes + i] = d[c];
    if constexpr ((DM::NP + DM::NF) > 0) {
        R::out(yy, pp, tt, pr, fl);
        if (prof)
#pragma unroll
            for (int c = 0; c < DM::NP; ++c) prof[c * nodes + i] = pr[c];
        if (field)
#pragma unroll
            for (int c = 0; c < DM::NF; ++c) field[c * nodes + i] = fl[c];
    }
}

// Transposed output Jacobians at `nodes` points: what csdl's reverse mode does with the profile / field outputs that
// VectorBased.integrate evaluates on the whole history (ozone/integrators/vector_based.py:33-49).
// vprof [NPROF][nodes], vfield [NFIELD][nodes] -> yb [D][nodes], pb [P][nodes], tb [S][nodes] (per node: the caller sums).
template <class R, typename T>
__global__ void __launch_bounds__(128) rhs_out_vjp_kernel(long long nodes, int th_shared, const T* __restrict__ y,
                                                          const T* __restrict__ p, const T* __restrict__ th,
                                                          const T* __restrict__ vprof, const T* __restrict__ vfield,
                                                          T* __restrict__ yb, T* __restrict__ pb, T* __restrict__ tb) {
    using DM = Dims<R>;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nodes) return;
    T yy[DM::D], pp[DM::P1], tt[DM::S1], vp[DM::NP1], vf[DM::NF1], gy[DM::D], gp[DM::P1], gt[DM::S1];
#pragma unroll
    for (int c = 0; c < DM::D; ++c) {
        yy[c] = y[c * nodes + i];
        gy[c] = T(0);
    }
#pragma unroll
    for (int c = 0; c < DM::P; ++c) pp[c] = p[c * nodes + i];
#pragma unroll
    for (int c = 0; c < DM::P1; ++c) gp[c] = T(0);
#pragma unroll
    for (int c = 0; c < DM::SP; ++c) tt[c] = th_shared ? th[c] : th[c * nodes + i];
#pragma unroll
    for (int c = 0; c < DM::S1; ++c) gt[c] = T(0);
#pragma unroll
    for (int c = 0; c < DM::NP; ++c) vp[c] = vprof ? vprof[c * nodes + i] : T(0);
#pragma unroll
    for (int c = 0; c < DM::NF; ++c) vf[c] = vfield ? vfield[c * nodes + i] : T(0);
    if constexpr ((DM::NP + DM::NF) > 0) R::out_vjp(yy, pp, tt, vp, vf, gy, gp, gt);
    if (yb)
#pragma unroll
        for (int c = 0; c < DM::D; ++c) yb[c * nodes + i] = gy[c];
    if (pb)
#pragma unroll
        for (int c = 0; c < DM::P; ++c) pb[c * nodes + i] = gp[c];
    if (tb)
#pragma unroll
        for (int c = 0; c < DM::SP; ++c) tb[c * nodes + i] = gt[c];
}

// Jacobians of f at blocks x nodes points: what an optimiser needs to write down the constraint Jacobian of the
// Collocation approach block by block (dc1/dY_k = h_k (A (x) I) blkdiag(J_y) - I, ...; ozone/integrators/collocation.py:48-57
// leaves this to csdl's reverse mode).  Layouts keep the trajectory index innermost, one block per (step, stage):
//   y [blocks][D][nodes], p [blocks][P][nodes]  ->  f [blocks][D][nodes], Jy [blocks][D][D][nodes] (d f_r / d y_c at [r][c]),
//   Jp [blocks][D][P][nodes]; th [S] or [S][nodes].  With blocks = n s and nodes = N these are the stage design variables
// of the Collocation approach as they lie in memory.
// COMPACT: only the entries the functor's patterns JP / JPP allow are written, in row-major order of the nonzeros
// (Jy [blocks][nnz_y][nodes], Jp [blocks][nnz_p][nodes]; oz_rhs_jac_pattern tells the caller which they are).  The lunar-ascent
// Jacobians have 9 of 49 and 3 of 21 entries: 19 instead of 77 values per node leave the chip.
// Stage mode (lerp.S > 0): block b is (step k, stage j) = (b / S, b % S) and `p` is the RAW parameter array at the grid
// points, [steps + 1][P][nodes] or, shared, [steps + 1][P]; the kernel interpolates it to the stage abscissa itself
// (ozone/collections.py:277-289), so no interpolated copy of the parameters is ever materialised.
template <typename T>
struct JacLerp {
    int S, p_shared;
    T C[kMaxS];
};

template <class R, typename T, bool COMPACT>
__global__ void __launch_bounds__(128) rhs_jac_kernel(long long nodes, int th_shared, const T* __restrict__ y,
                                                      const T* __restrict__ p, const T* __restrict__ th, T* __restrict__ f,
                                                      T* __restrict__ Jy, T* __restrict__ Jp, const JacLerp<T> lerp) {
    using DM = Dims<R>;
    constexpr int D = DM::D;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nodes) return;
    const long long b = blockIdx.y;
    y += b * D * nodes;
    T yy[D], pp[DM::P1], tt[DM::S1];
#pragma unroll
    for (int c = 0; c < D; ++c) yy[c] = y[c * nodes + i];
    if (lerp.S > 0) {
        const long long k = b / lerp.S;
        const T cj = lerp.C[b - k * lerp.S];
#pragma unroll
        for (int c = 0; c < DM::P; ++c) {
            const T lo = lerp.p_shared ? p[k * DM::P + c] : p[(k * DM::P + c) * nodes + i];
            const T hi = lerp.p_shared ? p[(k + 1) * DM::P + c] : p[((k + 1) * DM::P + c) * nodes + i];
            pp[c] = lo + cj * (hi - lo);
        }
    } else {
#pragma unroll
        for (int c = 0; c < DM::P; ++c) pp[c] = p[(b * DM::P + c) * nodes + i];
    }
#pragma unroll
    for (int c = 0; c < DM::SP; ++c) tt[c] = th_shared ? th[c] : th[c * nodes + i];
    if (f) {
        T d[D];
        R::f(yy, pp, tt, d);
#pragma unroll
        for (int c = 0; c < D; ++c) f[(b * D + c) * nodes + i] = d[c];
    }
    if (Jy) {
        T J[D * D];
        R::jac(yy, pp, tt, J);
        if constexpr (COMPACT) {
            constexpr int NNZ = jac_nnz_before<R>(D * D);
            static_for<0, D * D>([&](auto ec) {
                constexpr int e = decltype(ec)::value;
                if constexpr (jac_nz<R>(e / D, e % D)) Jy[(b * NNZ + jac_nnz_before<R>(e)) * nodes + i] = J[e];
            });
        } else {
#pragma unroll
            for (int c = 0; c < D * D; ++c) Jy[(b * D * D + c) * nodes + i] = J[c];
        }
    }
    if constexpr (DM::P > 0) {
        if (Jp) {
            constexpr int NNZP = jacp_nnz_before<R>(D * DM::P);
            static_for<0, D>([&](auto rc) {      // row r of J_p = J_p^T e_r
                constexpr int r = decltype(rc)::value;
                if constexpr (!COMPACT || jacp_nnz_before<R>((r + 1) * DM::P) > jacp_nnz_before<R>(r * DM::P)) {
                    T v[D], gy[D], gp[DM::P1], gt[DM::S1];
#pragma unroll
                    for (int c = 0; c < D; ++c) {
                        v[c] = c == r ? T(1) : T(0);
                        gy[c] = T(0);
                    }
#pragma unroll
                    for (int c = 0; c < DM::P1; ++c) gp[c] = T(0);
#pragma unroll
                    for (int c = 0; c < DM::S1; ++c) gt[c] = T(0);
                    R::vjp(yy, pp, tt, v, gy, gp, gt);
                    static_for<0, DM::P>([&](auto cc) {
                        constexpr int c = decltype(cc)::value;
                        if constexpr (!COMPACT)
                            Jp[((b * D + r) * DM::P + c) * nodes + i] = gp[c];
                        else if constexpr (jacp_nz<R>(r, c))
                            Jp[(b * NNZP + jacp_nnz_before<R>(r * DM::P + c)) * nodes + i] = gp[c];
                    });
                }
            });
        }
    }
}

}  // namespace oz
