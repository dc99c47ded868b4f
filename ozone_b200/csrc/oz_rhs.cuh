// Device right-hand sides registered behind func_wrapper.
//
// The reference calls a Python function f(ozone_vars) (ozone/func_wrapper.py:90-107) and lets
// csdl differentiate it.  Here each named problem is a device functor with the packed-vector
// contract below; everything is per trajectory, all arrays live in registers.
//
//   f      (y[D], p[P], th[S])            -> dy[D]
//   jac    (y, p, th)                     -> J[D*D]  row-major  J[i*D+j] = d f_i / d y_j   (implicit methods)
//   vjp    (y, p, th, v[D])               -> yb[D] += J_y^T v,  pb[P] += J_p^T v,  tb[S] += J_th^T v
//   out    (y, p, th)                     -> prof[NPROF], field[NFIELD]       (profile / field outputs)
//   out_vjp(y, p, th, vprof, vfield)      -> yb, pb, tb  +=  transposed output Jacobians
//   DQ                                    number of TRAILING states no right-hand side depends on (pure
//                                         quadratures, zero Jacobian columns): the implicit stage solve
//                                         eliminates them exactly and factors only the (D-DQ)-state system
//
// Packed orders follow the order in which the reference script adds states / parameters.
#pragma once
#include "oz_common.cuh"

namespace oz {

// examples/simple_examples/lotka_volterra.py:51-62.  y = (y, x); p = (a); th = (c).
struct RhsLotkaVolterra {
    static constexpr int ID = OZ_RHS_LOTKA_VOLTERRA, D = 2, P = 1, S = 1, NPROF = 1, NFIELD = 1;
    static constexpr int DQ = 0;
    // Written in factored form, y (a - x/2) and x (2y - c): the two factors are also the diagonal of the Jacobian,
    // so the reverse sweep's stage recompute and its vector-Jacobian product share them (4 + 8 operations per stage).
    template <typename T>
    OZ_D static void f(const T* y, const T* p, const T* th, T* dy) {
        dy[0] = y[0] * fma(T(-0.5), y[1], p[0]);
        dy[1] = y[1] * fma(T(2), y[0], -th[0]);
    }
    template <typename T>
    OZ_D static void jac(const T* y, const T* p, const T* th, T* J) {
        J[0] = fma(T(-0.5), y[1], p[0]);
        J[1] = -T(0.5) * y[0];
        J[2] = T(2) * y[1];
        J[3] = fma(T(2), y[0], -th[0]);
    }
    template <typename T>
    OZ_D static void vjp(const T* y, const T* p, const T* th, const T* v, T* yb, T* pb, T* tb) {
        const T vy = v[0] * y[0], vx = v[1] * y[1];   // also the parameter cotangents
        yb[0] = fma(T(2), vx, fma(v[0], fma(T(-0.5), y[1], p[0]), yb[0]));
        yb[1] = fma(T(-0.5), vy, fma(v[1], fma(T(2), y[0], -th[0]), yb[1]));
        pb[0] += vy;
        tb[0] -= vx;
    }
    template <typename T>
    OZ_D static void out(const T* y, const T*, const T*, T* prof, T* field) {
        prof[0] = y[1] + y[0];  // 'x_plus_y'
        field[0] = y[1];        // 'field_x'
    }
    template <typename T>
    OZ_D static void out_vjp(const T*, const T*, const T*, const T* vp, const T* vf, T* yb, T*, T*) {
        yb[0] += vp[0];
        yb[1] += vp[0] + vf[0];
    }
};

// ozone/paper_examples/van_der_pol/main_model.py:6-14.  y = (x0, x1, J); p = (u).
struct RhsVanDerPol {
    static constexpr int ID = OZ_RHS_VAN_DER_POL, D = 3, P = 1, S = 0, NPROF = 0, NFIELD = 0;
    static constexpr int DQ = 1;  // J is a pure quadrature: no right-hand side depends on it
    // which df_r / dy_c can be nonzero (oz_sparse_solve.cuh): x1' = x0 is linear and no derivative depends on J
    static constexpr unsigned char JP[9] = {1, 1, 0,
                                            1, 0, 0,
                                            1, 1, 0};
    template <typename T>
    OZ_D static void f(const T* y, const T* p, const T*, T* dy) {
        dy[0] = (T(1) - y[1] * y[1]) * y[0] - y[1] + p[0];
        dy[1] = y[0];
        dy[2] = y[0] * y[0] + y[1] * y[1] + p[0] * p[0];
    }
    template <typename T>
    OZ_D static void jac(const T* y, const T*, const T*, T* J) {
        J[0] = T(1) - y[1] * y[1];
        J[1] = -T(2) * y[0] * y[1] - T(1);
        J[2] = T(0);
        J[3] = T(1);
        J[4] = T(0);
        J[5] = T(0);
        J[6] = T(2) * y[0];
        J[7] = T(2) * y[1];
        J[8] = T(0);
    }
    template <typename T>
    OZ_D static void vjp(const T* y, const T* p, const T*, const T* v, T* yb, T* pb, T*) {
        yb[0] += v[0] * (T(1) - y[1] * y[1]) + v[1] + v[2] * T(2) * y[0];
        yb[1] += v[0] * (-T(2) * y[0] * y[1] - T(1)) + v[2] * T(2) * y[1];
        pb[0] += v[0] + v[2] * T(2) * p[0];
    }
    template <typename T>
    OZ_D static void out(const T*, const T*, const T*, T*, T*) {}
    template <typename T>
    OZ_D static void out_vjp(const T*, const T*, const T*, const T*, const T*, T*, T*, T*) {}
};

// ozone/paper_examples/ascent_system/main_model.py:6-21.  y = (rx, ry, rz, Vx, Vy, Vz, m); p = (ux, uy, uz).
struct RhsLunarAscent {
    static constexpr int ID = OZ_RHS_LUNAR_ASCENT, D = 7, P = 3, S = 0, NPROF = 0, NFIELD = 0;
    static constexpr int DQ = 0;
    static constexpr double kT = 1.8335634, kVex = 3049.87, kG0 = 1.625, kR0 = 1738100.0;
    static constexpr double kSqrtG0R0 = 1680.5988516002265;  // sqrt(g0 * R0) = sqrt(2824412.5)
    // possibly-nonzero entries of df/dy and df/dp (oz_sparse_solve.cuh; compact Jacobians for the optimiser adapter):
    // r' = V;  V' = -r + (T/m) u  (depends on r, m and u);  m' = const
    static constexpr unsigned char JP[49] = {0, 0, 0, 1, 0, 0, 0,   0, 0, 0, 0, 1, 0, 0,   0, 0, 0, 0, 0, 1, 0,
                                             1, 0, 0, 0, 0, 0, 1,   0, 1, 0, 0, 0, 0, 1,   0, 0, 1, 0, 0, 0, 1,
                                             0, 0, 0, 0, 0, 0, 0};
    static constexpr unsigned char JPP[21] = {0, 0, 0,  0, 0, 0,  0, 0, 0,  1, 0, 0,  0, 1, 0,  0, 0, 1,  0, 0, 0};
    template <typename T>
    OZ_D static void f(const T* y, const T* p, const T*, T* dy) {
        const T Tm = T(kT) / y[6];
        dy[0] = y[3];
        dy[1] = y[4];
        dy[2] = y[5];
        dy[3] = -y[0] + Tm * p[0];
        dy[4] = -y[1] + Tm * p[1];
        dy[5] = -y[2] + Tm * p[2];
        dy[6] = T(-(kT / kVex) * kSqrtG0R0);
    }
    template <typename T>
    OZ_D static void jac(const T* y, const T* p, const T*, T* J) {
#pragma unroll
        for (int i = 0; i < 49; ++i) J[i] = T(0);
        J[0 * 7 + 3] = T(1);
        J[1 * 7 + 4] = T(1);
        J[2 * 7 + 5] = T(1);
        const T q = -T(kT) / (y[6] * y[6]);
        J[3 * 7 + 0] = T(-1);
        J[4 * 7 + 1] = T(-1);
        J[5 * 7 + 2] = T(-1);
        J[3 * 7 + 6] = q * p[0];
        J[4 * 7 + 6] = q * p[1];
        J[5 * 7 + 6] = q * p[2];
    }
    template <typename T>
    OZ_D static void vjp(const T* y, const T* p, const T*, const T* v, T* yb, T* pb, T*) {
        const T Tm = T(kT) / y[6];
        yb[0] -= v[3];
        yb[1] -= v[4];
        yb[2] -= v[5];
        yb[3] += v[0];
        yb[4] += v[1];
        yb[5] += v[2];
        yb[6] -= (Tm / y[6]) * (v[3] * p[0] + v[4] * p[1] + v[5] * p[2]);
        pb[0] += v[3] * Tm;
        pb[1] += v[4] * Tm;
        pb[2] += v[5] * Tm;
    }
    template <typename T>
    OZ_D static void out(const T*, const T*, const T*, T*, T*) {}
    template <typename T>
    OZ_D static void out_vjp(const T*, const T*, const T*, const T*, const T*, T*, T*, T*) {}
};

// y' = th0*y + th1: test_1_simple/utils.py:6-7 (th = (-1, 0)); tutorial.ipynb (th = (0.5, -3)).
struct RhsAffine {
    static constexpr int ID = OZ_RHS_AFFINE, D = 1, P = 0, S = 2, NPROF = 0, NFIELD = 0;
    static constexpr int DQ = 0;
    template <typename T>
    OZ_D static void f(const T* y, const T*, const T* th, T* dy) { dy[0] = th[0] * y[0] + th[1]; }
    template <typename T>
    OZ_D static void jac(const T*, const T*, const T* th, T* J) { J[0] = th[0]; }
    template <typename T>
    OZ_D static void vjp(const T* y, const T*, const T* th, const T* v, T* yb, T*, T* tb) {
        yb[0] += v[0] * th[0];
        tb[0] += v[0] * y[0];
        tb[1] += v[0];
    }
    template <typename T>
    OZ_D static void out(const T*, const T*, const T*, T*, T*) {}
    template <typename T>
    OZ_D static void out_vjp(const T*, const T*, const T*, const T*, const T*, T*, T*, T*) {}
};

// ozone/tests/test_correctness/test_3_large/utils.py:23-30:  y' = -(y (dp + sp))^2.
struct RhsQuadDecay {
    static constexpr int ID = OZ_RHS_QUAD_DECAY, D = 1, P = 1, S = 1, NPROF = 0, NFIELD = 0;
    static constexpr int DQ = 0;
    template <typename T>
    OZ_D static void f(const T* y, const T* p, const T* th, T* dy) {
        const T w = y[0] * (p[0] + th[0]);
        dy[0] = -w * w;
    }
    template <typename T>
    OZ_D static void jac(const T* y, const T* p, const T* th, T* J) {
        const T q = p[0] + th[0];
        J[0] = -T(2) * y[0] * q * q;
    }
    template <typename T>
    OZ_D static void vjp(const T* y, const T* p, const T* th, const T* v, T* yb, T* pb, T* tb) {
        const T q = p[0] + th[0];
        yb[0] += v[0] * (-T(2) * y[0] * q * q);
        const T g = v[0] * (-T(2) * y[0] * y[0] * q);
        pb[0] += g;
        tb[0] += g;
    }
    template <typename T>
    OZ_D static void out(const T*, const T*, const T*, T*, T*) {}
    template <typename T>
    OZ_D static void out_vjp(const T*, const T*, const T*, const T*, const T*, T*, T*, T*) {}
};

// ozone/tests/test_correctness/test_2_general/utils.py:43-84.
// y = (y, x, z00, z01, z10, z11); p = (a, b, g, e00, e01, e10, e11, coeffs); th = (d).
// profile = ('z' -> z11, 'x' -> x + y^2/4, 'y' -> z*z01*z10 + e^2 + a + d (2x2)); field = ('y' -> c*y, 'z' -> c*z).
struct RhsGeneral3 {
    static constexpr int ID = OZ_RHS_GENERAL3, D = 6, P = 8, S = 1, NPROF = 6, NFIELD = 5;
    static constexpr int DQ = 0;
    template <typename T>
    OZ_D static void f(const T* y, const T* p, const T* th, T* dy) {
        dy[0] = p[0] * y[0] - p[1] * y[0] * y[1];
        dy[1] = p[2] * y[1] * y[0] - th[0] * y[1];
        const T common = y[0] * y[0] * T(0.5) + y[1] * T(0.5);
#pragma unroll
        for (int k = 0; k < 4; ++k) dy[2 + k] = -y[2 + k] / T(3) * y[3] + common + p[0] * p[3 + k];
    }
    template <typename T>
    OZ_D static void jac(const T* y, const T* p, const T* th, T* J) {
#pragma unroll
        for (int i = 0; i < 36; ++i) J[i] = T(0);
        J[0] = p[0] - p[1] * y[1];
        J[1] = -p[1] * y[0];
        J[6] = p[2] * y[1];
        J[7] = p[2] * y[0] - th[0];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            J[(2 + k) * 6 + 0] = y[0];
            J[(2 + k) * 6 + 1] = T(0.5);
            J[(2 + k) * 6 + 2 + k] += -y[3] / T(3);
            J[(2 + k) * 6 + 3] += -y[2 + k] / T(3);
        }
    }
    template <typename T>
    OZ_D static void vjp(const T* y, const T* p, const T* th, const T* v, T* yb, T* pb, T* tb) {
        T sv = T(0), sve = T(0), svz = T(0);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            sv += v[2 + k];
            sve += v[2 + k] * p[3 + k];
            svz += v[2 + k] * y[2 + k];
            yb[2 + k] -= v[2 + k] * y[3] / T(3);
            pb[3 + k] += v[2 + k] * p[0];
        }
        yb[3] -= svz / T(3);
        yb[0] += v[0] * (p[0] - p[1] * y[1]) + v[1] * p[2] * y[1] + sv * y[0];
        yb[1] += -v[0] * p[1] * y[0] + v[1] * (p[2] * y[0] - th[0]) + T(0.5) * sv;
        pb[0] += v[0] * y[0] + sve;
        pb[1] -= v[0] * y[0] * y[1];
        pb[2] += v[1] * y[1] * y[0];
        tb[0] -= v[1] * y[1];
    }
    template <typename T>
    OZ_D static void out(const T* y, const T* p, const T* th, T* prof, T* field) {
        prof[0] = y[5];
        prof[1] = y[1] + (y[0] / T(4)) * y[0];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            prof[2 + k] = y[2 + k] * y[3] * y[4] + p[3 + k] * p[3 + k] + p[0] + th[0];
            field[1 + k] = p[7] * y[2 + k];
        }
        field[0] = p[7] * y[0];
    }
    template <typename T>
    OZ_D static void out_vjp(const T* y, const T* p, const T*, const T* vp, const T* vf, T* yb, T* pb, T* tb) {
        yb[5] += vp[0];
        yb[1] += vp[1];
        yb[0] += vp[1] * y[0] * T(0.5);
        T sw = T(0), swz = T(0), sfz = T(0);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const T w = vp[2 + k];
            sw += w;
            swz += w * y[2 + k];
            sfz += vf[1 + k] * y[2 + k];
            yb[2 + k] += w * y[3] * y[4] + vf[1 + k] * p[7];
            pb[3 + k] += T(2) * p[3 + k] * w;
        }
        yb[3] += swz * y[4];
        yb[4] += swz * y[3];
        pb[0] += sw;
        tb[0] += sw;
        yb[0] += vf[0] * p[7];
        pb[7] += vf[0] * y[0] + sfz;
    }
};

}  // namespace oz
