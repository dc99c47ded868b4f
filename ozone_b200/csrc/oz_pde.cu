// The built-in grid model: nonlinear diffusion-reaction PDE (ozone/paper_examples/pde_control/pde_control.py:11-81).
//
//   x_j' = (x_{j-1} - 2 x_j + x_{j+1}) / dx^2 + beta_T (exp(-gamma/(1+x_j)) - exp(-gamma)) + beta_U (H_j u - x_j)
//   zero Dirichlet ghosts (:32-35), H_j = 1 for round(0.3 nx) <= j < round(0.7 nx) (:15-16)
//   field output J_space = (100/nx) sum_j x_j^2 + 20 u^2 with the rebound dx = 1/nx (:62-76)
//
// The kernels and the host-side sweep are in oz_pde_kernels.cuh, shared with grid callbacks traced from Python.
#include "oz_pde_kernels.cuh"

namespace oz {

template <typename T_>
struct DiffusionReaction {
    using T = T_;
    static constexpr int NCOEF = 0;
    static constexpr int P = 1;
    static constexpr bool WINDOW = true;
    T inv_dx2, beta_t, beta_u, gamma, bt_emg;
    long long h_lo, h_hi;  // global cell range where the control acts

    void configure(const oz_problem* pb, long long nx_global, T* field_scale, T* fb_scale) {
        const double dx = pb->rhs_cfg[0], bt = pb->rhs_cfg[1], bu = pb->rhs_cfg[2], g = pb->rhs_cfg[3];
        inv_dx2 = (T)(1.0 / (dx * dx));
        beta_t = (T)bt;
        beta_u = (T)bu;
        gamma = (T)g;
        bt_emg = (T)(bt * exp(-g));
        h_lo = (long long)pb->rhs_cfg[4];
        h_hi = (long long)pb->rhs_cfg[5];
        *field_scale = (T)(100.0 / (double)nx_global);
        *fb_scale = (T)(200.0 / (double)nx_global);
    }
    OZ_D T stage_u(T ulo, T du, T c) const { return beta_u * (ulo + c * du); }   // beta_U * u at the stage abscissa
    // x_j' and d(reaction + transfer)/dx from ONE exponential (the reverse step needs both at every stage)
    template <bool WITH_DIAG>
    OZ_D T f(T xl, T x, T xr, T mask, const T*, const T* bu_u, const T* etab, T& dg) const {
        const T bu_ui = bu_u[0];
        const T q = pde_rcp(T(1) + x);
        const T ex = beta_t * pde_exp(-gamma * q, etab);
        if constexpr (WITH_DIAG) dg = ex * gamma * q * q - beta_u;
        return (xl - T(2) * x + xr) * inv_dx2 + ex - bt_emg + (mask * bu_ui - beta_u * x);
    }
    OZ_D T jt(T gl, T g, T gr, T dg, const T*) const { return (gl - T(2) * g + gr) * inv_dx2 + dg * g; }   // the stencil is symmetric
    OZ_D T du(T mask, const T*, int) const { return mask; }
    OZ_D T du_scale() const { return beta_u; }
    OZ_D T q(T y, const T*) const { return y * y; }
    OZ_D T dq(T y, const T*) const { return y; }
    OZ_D T ucost(T own, const T* u) const { return own * T(20) * u[0] * u[0]; }
    OZ_D T ducost(T fb, const T* u, int) const { return T(40) * fb * u[0]; }
};

int pde_dispatch(const oz_problem* pb, int mode, size_t* ws_bytes, const EvalArgs* ev) {
    if (int rc = pde_check(pb, mode, ev)) return rc;
    return pb->dtype == OZ_F64 ? pde_by_stage<DiffusionReaction<double>>(pb, mode, ws_bytes)
                               : pde_by_stage<DiffusionReaction<float>>(pb, mode, ws_bytes);
}

}  // namespace oz
