// Per-thread small dense LU with partial pivoting, matrix held in shared memory.
//
// One thread owns one (M x M) system.  Element (r, c) of thread t lives at
//   base[(r*M + c) * stride + t]          (stride = threads per block)
// so that a warp touching the same (r, c) hits 32 consecutive words: conflict-free, no padding.
// Used by the implicit stage solve (Newton matrix I - h (A (x) I) blk(J), time_marching.py:235)
// and by its transposed (adjoint) solve.
#pragma once
#include "oz_common.cuh"

namespace oz {

template <typename T, int M>
struct SmemSystem {
    T* mat;    // already offset by threadIdx.x
    T* vec;    // already offset by threadIdx.x
    int* piv;  // already offset by threadIdx.x
    int stride;
    OZ_D T& m(int r, int c) { return mat[(r * M + c) * stride]; }
    OZ_D T& v(int r) { return vec[r * stride]; }
    OZ_D int& p(int r) { return piv[r * stride]; }

    static size_t bytes_per_thread() { return (size_t)M * M * sizeof(T) + (size_t)M * sizeof(T) + (size_t)M * sizeof(int); }

    OZ_D static SmemSystem carve(unsigned char* smem, int nthreads, int tid) {
        SmemSystem s;
        T* mat = reinterpret_cast<T*>(smem);
        T* vec = mat + (size_t)M * M * nthreads;
        int* piv = reinterpret_cast<int*>(vec + (size_t)M * nthreads);
        s.mat = mat + tid;
        s.vec = vec + tid;
        s.piv = piv + tid;
        s.stride = nthreads;
        return s;
    }

    // In-place P*Mat = L*U (unit lower L).  Returns false when a pivot is exactly zero / non-finite.
    OZ_D bool factor() {
        bool ok = true;
        for (int k = 0; k < M; ++k) {
            int pr = k;
            T best = fabs(m(k, k));
            for (int r = k + 1; r < M; ++r) {
                const T cand = fabs(m(r, k));
                if (cand > best) {
                    best = cand;
                    pr = r;
                }
            }
            p(k) = pr;
            if (pr != k) {
                for (int c = 0; c < M; ++c) {
                    const T t = m(k, c);
                    m(k, c) = m(pr, c);
                    m(pr, c) = t;
                }
            }
            const T piv_v = m(k, k);
            if (!(best > T(0)) || !isfinite(best)) {
                ok = false;
                continue;
            }
            const T inv = T(1) / piv_v;
            for (int r = k + 1; r < M; ++r) {
                const T l = m(r, k) * inv;
                m(r, k) = l;
                for (int c = k + 1; c < M; ++c) m(r, c) -= l * m(k, c);
            }
        }
        return ok;
    }

    // vec <- Mat^-1 vec
    OZ_D void solve() {
        for (int k = 0; k < M; ++k) {
            const int pr = p(k);
            if (pr != k) {
                const T t = v(k);
                v(k) = v(pr);
                v(pr) = t;
            }
        }
        for (int r = 1; r < M; ++r) {
            T acc = v(r);
            for (int c = 0; c < r; ++c) acc -= m(r, c) * v(c);
            v(r) = acc;
        }
        for (int r = M - 1; r >= 0; --r) {
            T acc = v(r);
            for (int c = r + 1; c < M; ++c) acc -= m(r, c) * v(c);
            v(r) = acc / m(r, r);
        }
    }

    // vec <- Mat^-T vec :  Mat^T = U^T L^T P  =>  x = P^T L^-T U^-T b
    OZ_D void solve_transposed() {
        for (int r = 0; r < M; ++r) {  // U^T is lower triangular
            T acc = v(r);
            for (int c = 0; c < r; ++c) acc -= m(c, r) * v(c);
            v(r) = acc / m(r, r);
        }
        for (int r = M - 2; r >= 0; --r) {  // L^T is unit upper triangular
            T acc = v(r);
            for (int c = r + 1; c < M; ++c) acc -= m(c, r) * v(c);
            v(r) = acc;
        }
        for (int k = M - 1; k >= 0; --k) {
            const int pr = p(k);
            if (pr != k) {
                const T t = v(k);
                v(k) = v(pr);
                v(pr) = t;
            }
        }
    }
};


// ---------------------------------------------------------------------------------------------
// Register-resident dense solve for small systems (M <= ~9): Gaussian elimination of the augmented
// system [a | b] with every index static, so the matrix never leaves the register file (the shared-
// memory LU above is bound by LDS/STS traffic: 3 shared accesses per DFMA).
//
// Threshold partial pivoting: rows are exchanged only when |a_kk| < 0.1 max_r |a_rk| (growth then stays
// bounded by 11^(M-1) in theory, ~O(10) in practice).  The Newton matrices I - h A (x) J almost never need
// an exchange, so the register path contains none: per column it only TESTS the diagonal (one compare per
// candidate row, no selects, no row copies -- an in-register exchange made the compiler pre-copy rows on
// the common path, 21 % of all issued instructions in the Gauss-Legendre forward sweep).  When a column
// fails the test the caller assembles the system again in local memory and solves it there with
// finish_with_pivoting(), a cold out-of-line routine with the same pivot rule.  The decision depends only on the thread's own
// data, so a trajectory's result does not depend on its warp neighbours.  On return b holds the solution;
// false = a zero / non-finite pivot was met.

// Reciprocal of a pivot: hardware seed r0 (rcp.approx, ~20 bits) corrected once to third order,
// r = r0 (1 + e + e^2) with e = 1 - x r0: the error is e^3 < 2^-60, i.e. full double precision after three dependent
// FMAs (two Newton steps need four) and without the special-case branches of an IEEE division (pivots here are O(1);
// a zero / non-finite pivot is flagged by the caller).
OZ_D double fast_rcp(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    const double e = fma(-x, r, 1.0);
    return fma(r, fma(e, e, e), r);
}
OZ_D float fast_rcp(float x) { return 1.0f / x; }

OZ_D int mag_bits(double x) { return __double2hiint(x) & 0x7fffffff; }
OZ_D int mag_bits(float x) { return __float_as_int(x) & 0x7fffffff; }

// Columns 0 .. k0-1 of the row-major M x M system are already eliminated (their diagonals hold reciprocals).
// Finishes the elimination from column k0 with threshold partial pivoting and back-substitutes.
template <typename T, int M>
__device__ __noinline__ bool finish_with_pivoting(T* a, T* b, int k0) {
    bool ok = true;
    for (int k = k0; k < M; ++k) {
        if (k < M - 1) {
            int pr = k;
            T best = fabs(a[k * M + k]);
            for (int r = k + 1; r < M; ++r) {
                const T cand = fabs(a[r * M + k]);
                if (cand > best) {
                    best = cand;
                    pr = r;
                }
            }
            if (pr != k && T(10) * fabs(a[k * M + k]) < best) {
                for (int c = k; c < M; ++c) {
                    const T t = a[k * M + c];
                    a[k * M + c] = a[pr * M + c];
                    a[pr * M + c] = t;
                }
                const T t = b[k];
                b[k] = b[pr];
                b[pr] = t;
            }
        }
        const T piv = a[k * M + k];
        if (!(fabs(piv) > T(0)) || !isfinite(piv)) ok = false;
        const T inv = fast_rcp(piv);
        a[k * M + k] = inv;
        for (int r = k + 1; r < M; ++r) {
            const T l = a[r * M + k] * inv;
            for (int c = k + 1; c < M; ++c) a[r * M + c] -= l * a[k * M + c];
            b[r] -= l * b[k];
        }
    }
    for (int r = M - 1; r >= 0; --r) {
        T acc = b[r];
        for (int c = r + 1; c < M; ++c) acc -= a[r * M + c] * b[c];
        b[r] = acc * a[r * M + r];
    }
    return ok;
}

// (column K is a template parameter: plain "#pragma unroll" on the column loop is not honoured by nvcc
// beyond M = 4, which turns every index dynamic and sends the matrix to local memory)
template <typename T, int M, int K>
struct RegElim {
    // bad: first column whose diagonal is below a tenth of an entry under it (elimination stops there), else -1
    OZ_D static void run(T (&a)[M][M], T (&b)[M], bool& ok, int& bad) {
        if constexpr (K < M - 1) {
            // magnitudes are compared through their leading 32 bits (monotone for non-negative IEEE numbers, NaN
            // and infinity rank highest): integer compares off the fp64 pipe; the threshold is a heuristic anyway
            const int lim = mag_bits(T(10) * fabs(a[K][K]));
            bool small = false;
#pragma unroll
            for (int r = K + 1; r < M; ++r) small = small || (mag_bits(a[r][K]) > lim);
            if (small) {
                bad = K;
                return;
            }
        }
        const T piv = a[K][K];
        if (!(fabs(piv) > T(0)) || !isfinite(piv)) ok = false;
        const T inv = fast_rcp(piv);
        a[K][K] = inv;
#pragma unroll
        for (int r = K + 1; r < M; ++r) {
            const T l = a[r][K] * inv;
#pragma unroll
            for (int c = K + 1; c < M; ++c) a[r][c] -= l * a[K][c];
            b[r] -= l * b[K];
        }
        if constexpr (K + 1 < M) RegElim<T, M, K + 1>::run(a, b, ok, bad);
    }
};

template <typename T, int M, int Rr>
struct RegBack {
    OZ_D static void run(T (&a)[M][M], T (&b)[M]) {
        // highest column first: the unknown found last (b[Rr + 1]) enters last, so everything else of this row's sum is
        // already done when it arrives and the dependent chain per row is one FMA and one multiplication
        T acc = b[Rr];
#pragma unroll
        for (int c = M - 1; c > Rr; --c) acc -= a[Rr][c] * b[c];
        b[Rr] = acc * a[Rr][Rr];
        if constexpr (Rr > 0) RegBack<T, M, Rr - 1>::run(a, b);
    }
};

// Returns 0 (solved, b = solution), 1 (a zero / non-finite pivot was met) or 2: some column failed the threshold
// test and nothing useful is left in a / b.  On 2 the caller builds the system again into row-major local arrays
// and calls finish_with_pivoting(ta, tb, 0).  (Rebuilding, rather than handing over the partly eliminated
// registers, keeps the cold path from needing any register state of the hot one: with a hand-over the compiler
// hoisted the row copies for it above the branch.)
template <typename T, int M>
OZ_D int reg_solve(T (&a)[M][M], T (&b)[M]) {
    bool ok = true;
    int bad = -1;
    RegElim<T, M, 0>::run(a, b, ok, bad);
    if (bad >= 0) return 2;
    RegBack<T, M, M - 1>::run(a, b);
    return ok ? 0 : 1;
}

}  // namespace oz
