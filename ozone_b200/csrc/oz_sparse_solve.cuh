// Structure-aware register elimination for the implicit stage systems.
//
// The Newton matrix of an implicit Runge-Kutta step, block (i, j) = delta_ij I - h a_ij J_f(Y_j)
// (ozone/integrators/time_marching.py:235), inherits every structural zero of the right-hand side's
// Jacobian: a state whose derivative does not depend on itself has a unit diagonal, a derivative that
// does not depend on state c leaves whole S x S blocks empty.  A functor may declare that pattern,
//
//     static constexpr unsigned char JP[D * D] = { ... };      // 0: df_r/dy_c is identically zero
//
// and this header turns it into straight-line code: the unknowns are ordered component by component
// (unit-diagonal components first), the fill-in of Gaussian elimination is worked out at COMPILE time
// (constexpr), and only entries that can be nonzero exist as registers or cost an instruction.  For
// Van der Pol under Gauss-Legendre 6 (x1' = x0 is linear) the 6 x 6 elimination collapses into a 3 x 3
// Schur complement: 27 + 9 products to form it, a dense 3 x 3 solve, 9 products to recover the rest.
// Functors without JP are treated as dense and get the same arithmetic as the dense register solve.
//
// Pivoting: same rule as reg_solve (oz_linalg.cuh).  Each column only TESTS its diagonal against the
// entries that can be nonzero below it; on failure the caller re-assembles densely and solves in the
// cold pivoting routine.  The decision depends on the thread's own data only.
#pragma once
#include <type_traits>
#include "oz_linalg.cuh"

namespace oz {

#ifndef OZ_SOLVE_CLOSED_TAIL   /* 1: a dense trailing block of at most 3 x 3 is solved in closed form (branch-free solves only).  Measured on
                                  Van der Pol GL6: 0.195 -> 0.200 ms (one reciprocal instead of three, but 33 products instead of
                                  ~20 on a pipe that is already 59 % busy): off. */
#define OZ_SOLVE_CLOSED_TAIL 0
#endif

template <int I>
struct IC {
    static constexpr int value = I;
};
// f(IC<B>{}), f(IC<B+1>{}), ... f(IC<E-1>{}): a loop whose index is a constant expression inside the body
template <int B, int E, class F>
OZ_D void static_for(F&& f) {
    if constexpr (B < E) {
        f(IC<B>{});
        static_for<B + 1, E>(f);
    }
}

template <class R, class = void>
struct HasJP : std::false_type {};
template <class R>
struct HasJP<R, std::void_t<decltype(R::JP)>> : std::true_type {};

// may df_r / dy_c be nonzero?
template <class R>
constexpr bool jac_nz(int r, int c) {
    if constexpr (HasJP<R>::value)
        return R::JP[r * R::D + c] != 0;
    else
        return true;
}

template <class R, class = void>
struct HasJPP : std::false_type {};
template <class R>
struct HasJPP<R, std::void_t<decltype(R::JPP)>> : std::true_type {};
// may df_r / dp_c be nonzero?  (optional functor member JPP[D * P], same convention as JP)
template <class R>
constexpr bool jacp_nz(int r, int c) {
    if constexpr (HasJPP<R>::value)
        return R::JPP[r * (R::P > 0 ? R::P : 1) + c] != 0;
    else
        return true;
}
// number of possibly-nonzero entries of df/dy (row-major) before entry e, and in total: the compact Jacobian layout
template <class R>
constexpr int jac_nnz_before(int e) {
    int cnt = 0;
    for (int q = 0; q < e; ++q) cnt += jac_nz<R>(q / R::D, q % R::D) ? 1 : 0;
    return cnt;
}
template <class R>
constexpr int jacp_nnz_before(int e) {
    int cnt = 0;
    for (int q = 0; q < e; ++q) cnt += jacp_nz<R>(q / (R::P > 0 ? R::P : 1), q % (R::P > 0 ? R::P : 1)) ? 1 : 0;
    return cnt;
}

template <int M>
struct Pat {           // 0: structurally zero, 1: exactly one, 2: general
    unsigned char v[M][M];
};

template <class R>
struct StageOrder {
    static constexpr int DA = (R::D - R::DQ) > 0 ? (R::D - R::DQ) : 1;
    int pos[DA];       // position of active component r in the elimination order
};
template <class R>
constexpr StageOrder<R> stage_order() {
    StageOrder<R> o{};
    constexpr int DA = R::D - R::DQ;
    int k = 0;
    for (int r = 0; r < DA; ++r)
        if (!jac_nz<R>(r, r)) o.pos[r] = k++;     // unit pivots first: they cost no reciprocal and no test
    for (int r = 0; r < DA; ++r)
        if (jac_nz<R>(r, r)) o.pos[r] = k++;
    return o;
}
// index of unknown (stage i, active component r) in the ordered system
template <class R, int S>
constexpr int uidx(int i, int r) {
    return stage_order<R>().pos[r] * S + i;
}

// pattern of the (transposed, if TR) Newton matrix in that order, assuming a full tableau
template <class R, int S, bool TR>
constexpr Pat<S*(R::D - R::DQ)> newton_pattern() {
    constexpr int DA = R::D - R::DQ;
    Pat<S * DA> p{};
    for (int i = 0; i < S; ++i)
        for (int r = 0; r < DA; ++r)
            for (int j = 0; j < S; ++j)
                for (int c = 0; c < DA; ++c) {
                    const int row = uidx<R, S>(i, r), col = uidx<R, S>(j, c);
                    const unsigned char t = jac_nz<R>(r, c) ? 2 : ((i == j && r == c) ? 1 : 0);
                    if (TR)
                        p.v[col][row] = t;
                    else
                        p.v[row][col] = t;
                }
    return p;
}
// pattern after columns 0 .. K-1 have been eliminated (entries below the diagonal of those columns are dead)
template <int M>
constexpr Pat<M> eliminated(Pat<M> p, int K) {
    for (int k = 0; k < K; ++k)
        for (int r = k + 1; r < M; ++r)
            if (p.v[r][k])
                for (int c = k + 1; c < M; ++c)
                    if (p.v[k][c]) p.v[r][c] = 2;
    return p;
}
// number of entries that ever hold a general value: what the register solve has to keep
template <class R, int S>
constexpr int newton_live_entries() {
    constexpr int M = S * (R::D - R::DQ);
    const Pat<M> p = eliminated<M>(newton_pattern<R, S, false>(), M);
    int cnt = 0;
    for (int r = 0; r < M; ++r)
        for (int c = 0; c < M; ++c) cnt += p.v[r][c] == 2;
    return cnt;
}

// First column K from which the remaining (M - K) x (M - K) block is dense and at most 3 x 3: that tail is solved in closed
// form (adjugate) instead of by elimination -- one reciprocal instead of three dependent ones, cofactors and right-hand
// side products in parallel.  M when there is no such tail.  (Van der Pol / Gauss-Legendre 6: K = 3, the Schur complement.)
template <class R, int S, bool TR>
constexpr int tail_start() {
    constexpr int M = S * (R::D - R::DQ);
    for (int K = 0; K < M; ++K) {
        if (M - K > 3 || M - K < 2) continue;
        const Pat<M> p = eliminated<M>(newton_pattern<R, S, TR>(), K);
        bool dense = true;
        for (int r = K; r < M; ++r)
            for (int c = K; c < M; ++c) dense = dense && p.v[r][c] == 2;
        if (dense) return K;
    }
    return M;
}

// x = A^-1 b for the dense tail rows / columns K .. M-1 (2 x 2 or 3 x 3), b overwritten.  `bad` is raised when |det| is
// tiny against the product of the row 1-norms (an upper bound of |det|): the caller then solves with row exchanges.
template <typename T, int M, int K>
OZ_D void solve_tail_closed(T (&a)[M][M], T (&b)[M], bool& ok, bool& bad) {
    if constexpr (M - K == 2) {
        const T a00 = a[K][K], a01 = a[K][K + 1], a10 = a[K + 1][K], a11 = a[K + 1][K + 1];
        const T det = a00 * a11 - a01 * a10;
        const T scale = (fabs(a00) + fabs(a01)) * (fabs(a10) + fabs(a11));
        bad = bad | !(fabs(det) > T(1e-8) * scale);
        ok = ok & (fabs(det) > T(0)) & (fabs(det) < CUDART_INF);
        const T inv = fast_rcp(det);
        const T x0 = a11 * b[K] - a01 * b[K + 1], x1 = a00 * b[K + 1] - a10 * b[K];
        b[K] = x0 * inv;
        b[K + 1] = x1 * inv;
    } else {
        const T a00 = a[K][K], a01 = a[K][K + 1], a02 = a[K][K + 2];
        const T a10 = a[K + 1][K], a11 = a[K + 1][K + 1], a12 = a[K + 1][K + 2];
        const T a20 = a[K + 2][K], a21 = a[K + 2][K + 1], a22 = a[K + 2][K + 2];
        // adjugate, row by row of the inverse: inv(A)[i][j] = cof[j][i] / det
        const T c00 = a11 * a22 - a12 * a21, c01 = a12 * a20 - a10 * a22, c02 = a10 * a21 - a11 * a20;
        const T c10 = a02 * a21 - a01 * a22, c11 = a00 * a22 - a02 * a20, c12 = a01 * a20 - a00 * a21;
        const T c20 = a01 * a12 - a02 * a11, c21 = a02 * a10 - a00 * a12, c22 = a00 * a11 - a01 * a10;
        const T det = a00 * c00 + a01 * c01 + a02 * c02;
        const T scale = (fabs(a00) + fabs(a01) + fabs(a02)) * (fabs(a10) + fabs(a11) + fabs(a12)) * (fabs(a20) + fabs(a21) + fabs(a22));
        bad = bad | !(fabs(det) > T(1e-8) * scale);
        ok = ok & (fabs(det) > T(0)) & (fabs(det) < CUDART_INF);
        const T inv = fast_rcp(det);
        const T b0 = b[K], b1 = b[K + 1], b2 = b[K + 2];
        b[K] = (c00 * b0 + c10 * b1 + c20 * b2) * inv;
        b[K + 1] = (c01 * b0 + c11 * b1 + c21 * b2) * inv;
        b[K + 2] = (c02 * b0 + c12 * b1 + c22 * b2) * inv;
    }
}

// EARLY: leave the elimination at the first column that fails its pivot test (one branch per tested column).
// !EARLY: a failed test only raises `bad`, the elimination runs on and the caller discards its result.  Without the
// branches the whole solve is one basic block, which lets the instruction scheduler overlap the columns -- this code is
// bound by the 8-cycle dependent-issue latency of the FP64 pipe (Van der Pol GL6 forward sweep: 0.220 -> 0.195 ms) -- at
// the price of more live registers: the 128-register reverse sweep is faster with the branches (0.114 vs 0.122 ms).
template <class R, int S, bool TR, typename T, int K, bool EARLY>
struct SparseElim {
    static constexpr int M = S * (R::D - R::DQ);
    static constexpr Pat<M> P = eliminated<M>(newton_pattern<R, S, TR>(), K);
    OZ_D static void run(T (&a)[M][M], T (&b)[M], bool& ok, bool& bad) {
        if constexpr (!EARLY && OZ_SOLVE_CLOSED_TAIL && K == tail_start<R, S, TR>()) {
            solve_tail_closed<T, M, K>(a, b, ok, bad);
            return;
        }
        constexpr int pk = P.v[K][K];
        static_assert(pk != 0, "structurally zero pivot");
        if constexpr (K < M - 1) {
            int lim;
            if constexpr (pk == 2)
                lim = mag_bits(T(10) * fabs(a[K][K]));
            else
                lim = mag_bits(T(10));
            static_for<K + 1, M>([&](auto rc) {
                constexpr int r = decltype(rc)::value;
                if constexpr (P.v[r][K] != 0) bad = bad | (mag_bits(a[r][K]) > lim);
            });
            if constexpr (EARLY) {
                if (bad) return;
            }
        }
        T inv = T(1);
        if constexpr (pk == 2) {
            const T piv = a[K][K];
            ok = ok & (fabs(piv) > T(0)) & (fabs(piv) < CUDART_INF);
            inv = fast_rcp(piv);
            a[K][K] = inv;
        }
        static_for<K + 1, M>([&](auto rc) {
            constexpr int r = decltype(rc)::value;
            if constexpr (P.v[r][K] != 0) {
                T l = a[r][K];
                if constexpr (pk == 2) l *= inv;
                static_for<K + 1, M>([&](auto cc) {
                    constexpr int c = decltype(cc)::value;
                    if constexpr (P.v[K][c] != 0) {
                        if constexpr (P.v[r][c] == 2)
                            a[r][c] -= l * a[K][c];
                        else if constexpr (P.v[r][c] == 1)
                            a[r][c] = fma(-l, a[K][c], T(1));
                        else
                            a[r][c] = -(l * a[K][c]);
                    }
                });
                b[r] -= l * b[K];
            }
        });
        if constexpr (K + 1 < M) SparseElim<R, S, TR, T, K + 1, EARLY>::run(a, b, ok, bad);
    }
};

template <class R, int S, bool TR, typename T, int Rr, bool EARLY>
struct SparseBack {
    static constexpr int M = S * (R::D - R::DQ);
    static constexpr Pat<M> P = eliminated<M>(newton_pattern<R, S, TR>(), M);
    OZ_D static void run(T (&a)[M][M], T (&b)[M]) {
        if constexpr (!EARLY && OZ_SOLVE_CLOSED_TAIL && Rr >= tail_start<R, S, TR>()) {   // already solved in closed form
            if constexpr (Rr > 0) SparseBack<R, S, TR, T, Rr - 1, EARLY>::run(a, b);
            return;
        }
        // highest column first (see RegBack): the unknown found last enters last
        T acc = b[Rr];
        static_for<0, M - 1 - Rr>([&](auto kc) {
            constexpr int c = M - 1 - decltype(kc)::value;
            if constexpr (P.v[Rr][c] != 0) acc -= a[Rr][c] * b[c];
        });
        if constexpr (P.v[Rr][Rr] == 2)
            b[Rr] = acc * a[Rr][Rr];
        else
            b[Rr] = acc;
        if constexpr (Rr > 0) SparseBack<R, S, TR, T, Rr - 1, EARLY>::run(a, b);
    }
};

// Entries of `a` whose pattern is 0 or 1 are never read: the caller stores only the general ones (newton_pattern == 2).
// Returns 0 (solved, b = solution in uidx order), 1 (zero / non-finite pivot) or 2 (a column failed the threshold
// test: assemble densely and use finish_with_pivoting).
template <class R, int S, bool TR, typename T, bool EARLY = false>
OZ_D int sparse_solve(T (&a)[S * (R::D - R::DQ)][S * (R::D - R::DQ)], T (&b)[S * (R::D - R::DQ)]) {
    constexpr int M = S * (R::D - R::DQ);
    bool ok = true, bad = false;
    SparseElim<R, S, TR, T, 0, EARLY>::run(a, b, ok, bad);
    if constexpr (EARLY) {
        if (bad) return 2;
    }
    SparseBack<R, S, TR, T, M - 1, EARLY>::run(a, b);
    return bad ? 2 : (ok ? 0 : 1);
}

}  // namespace oz
