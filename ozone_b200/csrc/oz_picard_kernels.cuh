// Picard iteration: relax every (step, stage) node of a trajectory at once.
//
// Reference: ozone/integrators/picard_iteration.py:9-65.  Unknown Ybar (n, s, d).  One sweep:
//     hF      = h_k f(Ybar_ki, p_ki)                                           (:43-44)
//     S       = (I - shift (x) V)^-1 (B hF + y0vec)                            (:17-19, :46)
//     stages  = A hF + U S ;  residual = Ybar - stages ;  Ybar <- Ybar - residual   (:46-52)
// For every registered method V = [[1]], so (I - shift (x) V)^-1 is the lower-triangular all-ones
// block matrix: S is a PREFIX SUM over time.  The reference materialises that inverse with
// scipy.sparse.linalg.inv (O(nt^2 d) memory); here it is a warp scan.
//
// Mapping: one WARP per trajectory, the time axis across the 32 lanes (lane l owns the contiguous
// steps [l*c, (l+1)*c), c = ceil(n/32)).  The iterate, A.hF and the per-step increments live in
// shared memory in a lane-private layout (element j of thread t at buf[j*blockDim.x + t]:
// conflict-free), so all sweeps stay on-chip; HBM sees the inputs once and the solved history once.
// States are relaxed block by block (block Gauss-Seidel over the states in insertion order, as the
// csdl GaussSeidel solver does, :35); one block = Jacobi.
#pragma once
#include "oz_common.cuh"
#include "oz_tm_kernels.cuh"

namespace oz {

constexpr int kMaxBlocks = 8;

template <typename T>
struct PicardArgs {
    Tableau<T> tab;
    long long N;
    int n, chunk;
    int h_shared, p_shared, th_shared;
    int nblocks;
    int blk_lo[kMaxBlocks + 1];
    T tol;
    int max_sweeps;
    const T *y0, *h, *p, *th, *guess;
    T *hist, *yfinal, *prof, *field;
    int *status, *iters;
};

// C > 0: the lane's `chunk` of steps is the compile-time constant C and A.hF stays in REGISTERS (half the shared
// memory per trajectory -> twice the resident warps, no LDS/STS for it); C = 0: any chunk, A.hF in shared memory.
#ifndef OZ_PICARD_GS_MAX_D   /* compile-time per-component Gauss-Seidel blocks for right-hand sides with up to this many states */
#define OZ_PICARD_GS_MAX_D 3
#endif

template <class R, int S>
inline size_t picard_elems_per_lane(int chunk, bool af_in_regs) {
    return (size_t)(af_in_regs ? 1 : 2) * chunk * S * R::D + (size_t)chunk * R::D +
           (size_t)(chunk + 1) * AtLeast1<R::P>::value + chunk;
}

// GS = 1: every packed component is its own Gauss-Seidel block (scalar states, the usual case): the block loop is
// unrolled with compile-time bounds, so inside block b only component b of f, of A.hF and of the running sum is
// computed at all (round 1 evaluated everything for every block and threw the other components away).
// GS = 0: blocks come from the argument block; the products and scans of foreign components are skipped at run time.
template <class R, typename T, int S, int C, int GS = 0>
__global__ void __launch_bounds__(C > 0 ? 512 : 256) picard_kernel(const PicardArgs<T> a) {
    using DM = Dims<R>;
    constexpr int D = DM::D, P = DM::P;
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, stride = blockDim.x;
    const long long traj = (long long)blockIdx.x * (blockDim.x >> 5) + (tid >> 5);
    if (traj >= a.N) return;  // whole warp leaves together; only warp-level sync below
    const long long N = a.N, i = traj;
    const int n = a.n, c = C > 0 ? C : a.chunk;
    const int k0 = lane * c;
    const Tableau<T>& tab = a.tab;

    T* Yb = reinterpret_cast<T*>(smem) + tid;     // [c][S][D]
    T* AF = Yb + (size_t)c * S * D * stride;      // [c][S][D]  (C = 0 only)
    T* inc = AF + (C > 0 ? (size_t)0 : (size_t)c * S * D * stride);   // [c][D]   lane-local inclusive prefix of h sum_i b_i F_i
    T AFr[C > 0 ? C : 1][S][D];
    auto af_set = [&](int q, int s, int cc, T v) {
        if constexpr (C > 0) AFr[q][s][cc] = v; else AF[(size_t)((q * S + s) * D + cc) * stride] = v;
    };
    auto af_get = [&](int q, int s, int cc) -> T {
        if constexpr (C > 0) return AFr[q][s][cc]; else return AF[(size_t)((q * S + s) * D + cc) * stride];
    };
    T* pp = inc + (size_t)c * D * stride;         // [c+1][P] raw parameters at this lane's grid points
    T* hh = pp + (size_t)(c + 1) * DM::P1 * stride;  // [c]
#define LP(buf, j) (buf)[(size_t)(j) * stride]

    T y0[D], th[DM::S1];
#pragma unroll
    for (int cc = 0; cc < D; ++cc) y0[cc] = __ldg(a.y0 + cc * N + i);
#pragma unroll
    for (int cc = 0; cc < DM::SP; ++cc) th[cc] = ld_in(a.th, cc, N, i, a.th_shared);
    for (int q = 0; q <= c; ++q) {
        const int k = min(k0 + q, n);
#pragma unroll
        for (int cc = 0; cc < P; ++cc) LP(pp, q * P + cc) = ld_in(a.p, (long long)k * P + cc, N, i, a.p_shared);
    }
    for (int q = 0; q < c; ++q) {
        const int k = k0 + q;
        if (k >= n) break;
        LP(hh, q) = ld_in(a.h, k, N, i, a.h_shared);
        // initial iterate: user guess repeated per stage, else the broadcast IC (collections.py:125-133)
#pragma unroll
        for (int cc = 0; cc < D; ++cc) {
            const T g = a.guess ? __ldg(a.guess + ((long long)k * D + cc) * N + i) : y0[cc];
#pragma unroll
            for (int s = 0; s < S; ++s) LP(Yb, (q * S + s) * D + cc) = g;
        }
    }

    int sweeps = 0, status = 0;
    T off[D];  // exclusive warp prefix of the lane totals (valid after the last scan)
#pragma unroll
    for (int cc = 0; cc < D; ++cc) off[cc] = T(0);
    // One relaxation of the components [lo, hi).  B >= 0: lo = B, hi = B + 1 are compile-time constants.
    // Returns this block's residual 2-norm squared and adds the squares of the new iterate to yn2.
    auto relax_block = [&](auto btag, int lo_rt, int hi_rt, T& yn2) -> T {
        constexpr int B = decltype(btag)::value;
        const int lo = B >= 0 ? B : lo_rt, hi = B >= 0 ? B + 1 : hi_rt;
        // ---- pass 1: f at every node of this lane, A.hF and the running B.hF sum
        T run[D];
#pragma unroll
        for (int cc = 0; cc < D; ++cc) run[cc] = T(0);
#pragma unroll
        for (int q = 0; q < c; ++q) {
            if (k0 + q >= n) continue;
            T Y[S][D], F[S][D], ps[S][DM::P1], plo[DM::P1], phi[DM::P1];
            const T h = LP(hh, q);
#pragma unroll
            for (int cc = 0; cc < P; ++cc) {
                plo[cc] = LP(pp, q * P + cc);
                phi[cc] = LP(pp, (q + 1) * P + cc);
            }
            stage_params<R, T, S>(tab, plo, phi, ps);
#pragma unroll
            for (int s = 0; s < S; ++s) {
#pragma unroll
                for (int cc = 0; cc < D; ++cc) Y[s][cc] = LP(Yb, (q * S + s) * D + cc);
                R::f(Y[s], ps[s], th, F[s]);
            }
#pragma unroll
            for (int cc = 0; cc < D; ++cc) {
                if (cc < lo || cc >= hi) continue;
                T accb = T(0);
#pragma unroll
                for (int s = 0; s < S; ++s) {
                    accb += tab.B[s] * F[s][cc];
                    T acca = T(0);
#pragma unroll
                    for (int j = 0; j < S; ++j) acca += tab.a(s, j) * F[j][cc];
                    af_set(q, s, cc, h * acca);
                }
                run[cc] += h * accb;
                LP(inc, q * D + cc) = run[cc];
            }
        }
        // ---- exclusive scan of the lane totals across the warp (the (I - V shift)^-1 product)
#pragma unroll
        for (int cc = 0; cc < D; ++cc) {
            if (cc < lo || cc >= hi) continue;
            T v = run[cc];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const T u = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += u;
            }
            off[cc] = v - run[cc];
        }
        // ---- pass 2: stages = A.hF + S_k ; residual and update for the components of this block
        T rn2 = T(0);
#pragma unroll
        for (int q = 0; q < c; ++q) {
            if (k0 + q >= n) continue;
#pragma unroll
            for (int cc = 0; cc < D; ++cc) {
                if (cc < lo || cc >= hi) continue;
                const T Sk = y0[cc] + off[cc] + (q > 0 ? LP(inc, (q - 1) * D + cc) : T(0));
#pragma unroll
                for (int s = 0; s < S; ++s) {
                    const T ynew = af_get(q, s, cc) + Sk;
                    const T r = LP(Yb, (q * S + s) * D + cc) - ynew;
                    rn2 += r * r;
                    yn2 += ynew * ynew;
                    LP(Yb, (q * S + s) * D + cc) = ynew;
                }
            }
        }
        return rn2;
    };
    // Round-off floor of the residual norm.  S_k is a running sum over up to n steps, so two successive iterates that are
    // both correct to working precision still differ by ~eps sqrt(n) |Y| per entry.  A sweep whose residual is below
    // 8 eps sqrt(n) ||Y||_2 and no longer halves has converged as far as this precision allows (fp32, n = 100, 800
    // unknowns of size ~2: floor ~3e-4, above the usual tolerance 1e-4; fp64: ~6e-13, below every tolerance in use).
    const T eps = sizeof(T) == 8 ? T(1.1102230246251565e-16) : T(5.9604645e-8);
    const T floor_scale = T(8) * eps * sqrt((T)n);
    T prev = CUDART_INF;
    for (;;) {
        T rmax = T(0), yn2 = T(0);
        if constexpr (GS == 1) {
            static_for<0, D>([&](auto bc) {
                T rn2 = warp_sum(relax_block(bc, 0, 0, yn2));
                rmax = fmax(rmax, sqrt(rn2));
                if (!(rn2 == rn2)) rmax = CUDART_INF;
            });
        } else {
            for (int b = 0; b < a.nblocks; ++b) {
                T rn2 = warp_sum(relax_block(IC<-1>{}, a.blk_lo[b], a.blk_lo[b + 1], yn2));
                rmax = fmax(rmax, sqrt(rn2));
                if (!(rn2 == rn2)) rmax = CUDART_INF;
            }
        }
        if (rmax < a.tol) break;
        if (!(rmax < CUDART_INF)) {
            status |= OZ_STATUS_NONFINITE;
            break;
        }
        if (rmax > T(0.5) * prev && rmax <= floor_scale * sqrt(warp_sum(yn2))) break;   // stalled on the round-off floor
        prev = rmax;
        if (sweeps >= a.max_sweeps) {
            status |= OZ_STATUS_NOT_CONVERGED;
            break;
        }
        ++sweeps;
    }
    // With several blocks the history must come from one consistent evaluation of f at the final
    // iterate (picard_iteration.py:56-65); the last block's pass 1 used iterates updated by earlier
    // blocks, which at convergence differ from the final ones by less than the tolerance.  Re-evaluate.
    if (a.nblocks > 1) {
        T run[D];
#pragma unroll
        for (int cc = 0; cc < D; ++cc) run[cc] = T(0);
        for (int q = 0; q < c; ++q) {
            if (k0 + q >= n) break;
            T Y[D], F[D], ps[S][DM::P1], plo[DM::P1], phi[DM::P1], accb[D];
            const T h = LP(hh, q);
#pragma unroll
            for (int cc = 0; cc < P; ++cc) {
                plo[cc] = LP(pp, q * P + cc);
                phi[cc] = LP(pp, (q + 1) * P + cc);
            }
            stage_params<R, T, S>(tab, plo, phi, ps);
#pragma unroll
            for (int cc = 0; cc < D; ++cc) accb[cc] = T(0);
#pragma unroll
            for (int s = 0; s < S; ++s) {
#pragma unroll
                for (int cc = 0; cc < D; ++cc) Y[cc] = LP(Yb, (q * S + s) * D + cc);
                R::f(Y, ps[s], th, F);
#pragma unroll
                for (int cc = 0; cc < D; ++cc) accb[cc] += tab.B[s] * F[cc];
            }
#pragma unroll
            for (int cc = 0; cc < D; ++cc) {
                run[cc] += h * accb[cc];
                LP(inc, q * D + cc) = run[cc];
            }
        }
#pragma unroll
        for (int cc = 0; cc < D; ++cc) {
            T v = run[cc];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const T u = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += u;
            }
            off[cc] = v - run[cc];
        }
    }

    // ---- outputs: history = S, profile/field on the solved history with the raw parameters
    constexpr bool HAS_OUT = (DM::NP + DM::NF) > 0;
    T facc[DM::NF1], prof[DM::NP1], fld[DM::NF1], pk[DM::P1], yk[D];
#pragma unroll
    for (int cc = 0; cc < DM::NF1; ++cc) facc[cc] = T(0);
    if (lane == 0) {
        if (a.hist)
#pragma unroll
            for (int cc = 0; cc < D; ++cc) a.hist[cc * N + i] = y0[cc];
        if constexpr (HAS_OUT) {
#pragma unroll
            for (int cc = 0; cc < P; ++cc) pk[cc] = LP(pp, cc);
            R::out(y0, pk, th, prof, facc);
            if (a.prof)
#pragma unroll
                for (int cc = 0; cc < DM::NP; ++cc) a.prof[cc * N + i] = prof[cc];
        }
    }
    for (int q = 0; q < c; ++q) {
        const int k = k0 + q;
        if (k >= n) break;
#pragma unroll
        for (int cc = 0; cc < D; ++cc) yk[cc] = y0[cc] + off[cc] + LP(inc, q * D + cc);  // y_{k+1}
        if (a.hist)
#pragma unroll
            for (int cc = 0; cc < D; ++cc) a.hist[((long long)(k + 1) * D + cc) * N + i] = yk[cc];
        if (k == n - 1 && a.yfinal)
#pragma unroll
            for (int cc = 0; cc < D; ++cc) a.yfinal[cc * N + i] = yk[cc];
        if constexpr (HAS_OUT) {
#pragma unroll
            for (int cc = 0; cc < P; ++cc) pk[cc] = LP(pp, (q + 1) * P + cc);
            R::out(yk, pk, th, prof, fld);
            if (a.prof)
#pragma unroll
                for (int cc = 0; cc < DM::NP; ++cc) a.prof[((long long)(k + 1) * DM::NP + cc) * N + i] = prof[cc];
#pragma unroll
            for (int cc = 0; cc < DM::NF; ++cc) facc[cc] += fld[cc];
        }
    }
    if constexpr (DM::NF > 0) {
#pragma unroll
        for (int cc = 0; cc < DM::NF; ++cc) {
            const T s = warp_sum(facc[cc]);
            if (lane == 0 && a.field) a.field[cc * N + i] = s;
        }
    }
    if (lane == 0) {
        if (a.status) a.status[i] = status;
        if (a.iters) a.iters[i] = sweeps;
    }
#undef LP
}

template <class R, typename T, int S>
int run_picard(const oz_problem* pb, const void* guess, double tol, int max_sweeps, const int* blk_lo, int nblocks) {
    PicardArgs<T> a;
    a.tab = make_tableau<T>(pb);
    a.N = pb->N;
    a.n = pb->n;
    a.chunk = (pb->n + 31) / 32;
    a.h_shared = pb->h_shared;
    a.p_shared = pb->p_shared;
    a.th_shared = pb->th_shared;
    a.nblocks = nblocks;
    for (int b = 0; b <= nblocks; ++b) a.blk_lo[b] = blk_lo[b];
    a.tol = (T)tol;
    a.max_sweeps = max_sweeps;
    a.y0 = (const T*)pb->y0;
    a.h = (const T*)pb->h;
    a.p = (const T*)pb->p;
    a.th = (const T*)pb->th;
    a.guess = (const T*)guess;
    a.hist = (T*)pb->hist;
    a.yfinal = (T*)pb->yfinal;
    a.prof = (T*)pb->prof;
    a.field = (T*)pb->field;
    a.status = pb->status;
    a.iters = pb->iters;
    // register variant when the lane's chunk is small and C * S * D doubles fit the register file comfortably
    const int creg = a.chunk <= 1 ? 1 : (a.chunk <= 2 ? 2 : (a.chunk <= 4 ? 4 : 0));
    const bool regs = creg > 0 && (size_t)creg * S * R::D * sizeof(T) <= 256;
    if (regs) a.chunk = creg;
    const size_t per_lane = picard_elems_per_lane<R, S>(a.chunk, regs) * sizeof(T);
    int warps = regs ? 16 : 8;
    while (warps > 1 && per_lane * 32 * warps > kSmemBudget) warps >>= 1;
    const size_t smem = per_lane * 32 * warps;
    OZ_REQUIRE(smem <= 227 * 1024,
               "Picard iterate of one trajectory (%zu bytes) does not fit in shared memory; use TimeMarching", smem);
    const long long nblk = (pb->N + warps - 1) / warps;
    OZ_REQUIRE(nblk < (1ll << 31), "batch too large for one launch");
    auto launch = [&](auto kern) -> int {
        if (smem > 48 * 1024) OZ_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)nblk, warps * 32, smem, (cudaStream_t)pb->stream>>>(a);
        OZ_CUDA_CHECK(cudaGetLastError());
        return 0;
    };
    // scalar states relaxed one by one (the reference's block Gauss-Seidel over scalar states): compile-time blocks
    bool single = nblocks == R::D && R::D >= 2;
    for (int b = 0; b <= nblocks && single; ++b) single = blk_lo[b] == b;
    if constexpr (R::D >= 2 && R::D <= OZ_PICARD_GS_MAX_D) {
        if (single) {
            if (regs && creg == 1) return launch(picard_kernel<R, T, S, 1, 1>);
            if (regs && creg == 2) return launch(picard_kernel<R, T, S, 2, 1>);
            if (regs && creg == 4) return launch(picard_kernel<R, T, S, 4, 1>);
            return launch(picard_kernel<R, T, S, 0, 1>);
        }
    }
    if (regs && creg == 1) return launch(picard_kernel<R, T, S, 1, 0>);
    if (regs && creg == 2) return launch(picard_kernel<R, T, S, 2, 0>);
    if (regs && creg == 4) return launch(picard_kernel<R, T, S, 4, 0>);
    return launch(picard_kernel<R, T, S, 0, 0>);
}

template <class R>
int dispatch_picard(const oz_problem* pb, const void* guess, double tol, int max_sweeps, const int* blk_lo, int nblocks) {
    OZ_REQUIRE(nblocks >= 1 && nblocks <= kMaxBlocks, "between 1 and %d state blocks are supported", kMaxBlocks);
    OZ_REQUIRE(blk_lo[0] == 0 && blk_lo[nblocks] == R::D, "state blocks must cover the %d packed components", R::D);
#define OZ_PICARD_CASE(SS)                                                                                  \
    if constexpr (OZ_WANT_S(SS) && OZ_WANT_KIND(1)) {                                                       \
        if (pb->num_stages == SS) {                                                                         \
            if constexpr (OZ_WANT_F64(true))                                                                \
                if (pb->dtype == OZ_F64) return run_picard<R, double, SS>(pb, guess, tol, max_sweeps, blk_lo, nblocks); \
            if constexpr (OZ_WANT_F64(false))                                                               \
                if (pb->dtype == OZ_F32) return run_picard<R, float, SS>(pb, guess, tol, max_sweeps, blk_lo, nblocks);  \
        }                                                                                                   \
    }
    OZ_PICARD_CASE(1)
    OZ_PICARD_CASE(2)
    OZ_PICARD_CASE(3)
    OZ_PICARD_CASE(4)
    OZ_PICARD_CASE(7)
#undef OZ_PICARD_CASE
    set_error("no Picard kernel for %d stages / dtype %d in this build", pb->num_stages, pb->dtype);
    return -1;
}

}  // namespace oz
