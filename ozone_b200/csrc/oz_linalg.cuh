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

}  // namespace oz
