"""Oracle right-hand sides (TEST INFRASTRUCTURE, see oracle/__init__.py).

Each class restates one user RHS of the reference in numpy.  States, dynamic parameters and
static parameters are *packed* along axis 0 in the order the reference script adds them;
every trailing axis is a batch axis (so shared inputs broadcast).  Everything is written with
operations that accept complex inputs so that ``jac``/``vjp`` default to complex-step
differentiation (exact to round-off, no hand-derived Jacobians to get wrong in the checker).
"""
import numpy as np

_EPS = 1e-30


class Rhs:
    name = ''
    D = P = S = 0            # packed sizes: state, dynamic parameter, static parameter
    NPROF = NFIELD = 0       # packed sizes of profile / field outputs

    def f(self, y, p, th):
        raise NotImplementedError

    def out(self, y, p, th):
        """-> (profile (NPROF, ...), field (NFIELD, ...)) evaluated at a grid point."""
        shp = np.broadcast_shapes(y.shape[1:], p.shape[1:] if self.P else (), th.shape[1:] if self.S else ())
        return np.zeros((0,) + shp), np.zeros((0,) + shp)

    # ---- derivatives by complex step (batch-vectorised) -------------------------------
    def _cs(self, fun, args, which):
        """Jacobian of fun(*args) (axis 0 = outputs) w.r.t. args[which] (axis 0 = inputs):
        result (n_out, n_in, ...)."""
        x = args[which]
        cols = []
        for j in range(x.shape[0]):
            xp = x.astype(complex)
            xp[j] = xp[j] + 1j * _EPS
            a = list(args)
            a[which] = xp
            cols.append(np.imag(fun(*a)) / _EPS)
        if not cols:
            base = fun(*args)
            return np.zeros((base.shape[0], 0) + base.shape[1:])
        cols = np.broadcast_arrays(*cols)
        return np.stack(cols, axis=1)

    def jac(self, y, p, th):
        """df/dy, shape (D, D, ...)."""
        return self._cs(self.f, (y, p, th), 0)

    def vjp(self, y, p, th, v):
        """(J_y^T v, J_p^T v, J_th^T v), each shaped like the (broadcast) argument."""
        args = (y, p, th)
        res = []
        for k in range(3):
            J = self._cs(self.f, args, k)
            res.append(np.einsum('i...,ij...->j...', v, J))
        return tuple(res)

    def out_vjp(self, y, p, th, vprof, vfield):
        args = (y, p, th)
        res = []
        for k in range(3):
            tot = 0.0
            for sel, v in ((0, vprof), (1, vfield)):
                J = self._cs(lambda *a: self.out(*a)[sel], args, k)
                tot = tot + np.einsum('i...,ij...->j...', v, J)
            res.append(tot)
        return tuple(res)


class LotkaVolterra(Rhs):
    """examples/simple_examples/lotka_volterra.py:51-62.  States (y, x); dynamic a(t); static c."""
    name = 'lotka_volterra'
    D, P, S, NPROF, NFIELD = 2, 1, 1, 1, 1

    def f(self, y, p, th):
        yy, xx = y[0], y[1]
        a, c = p[0], th[0]
        return np.stack(np.broadcast_arrays(a * yy - 0.5 * yy * xx, 2.0 * xx * yy - c * xx))

    def out(self, y, p, th):
        yy, xx = y[0], y[1]
        return (yy + xx)[None], (xx + 0 * yy)[None]      # profile x_plus_y, field field_x


class VanDerPol(Rhs):
    """ozone/paper_examples/van_der_pol/main_model.py:6-14.  States (x0, x1, J); dynamic u."""
    name = 'van_der_pol'
    D, P, S = 3, 1, 0

    def f(self, y, p, th):
        x0, x1, u = y[0], y[1], p[0]
        return np.stack(np.broadcast_arrays((1.0 - x1 ** 2) * x0 - x1 + u, x0 * 1.0 + 0 * u,
                                            x0 ** 2 + x1 ** 2 + u ** 2))


class LunarAscent(Rhs):
    """ozone/paper_examples/ascent_system/main_model.py:6-21.
    States (rx, ry, rz, Vx, Vy, Vz, m); dynamic (ux, uy, uz)."""
    name = 'lunar_ascent'
    D, P, S = 7, 3, 0
    g0, R0, vex, T = 1.625, 1738100, 3049.87, 1.8335634

    def f(self, y, p, th):
        Tm = self.T / y[6]
        mdot = -(self.T / self.vex) * np.sqrt(self.g0 * self.R0)
        rows = [y[3], y[4], y[5],
                -y[0] + Tm * p[0], -y[1] + Tm * p[1], -y[2] + Tm * p[2],
                mdot + 0 * y[6]]
        return np.stack(np.broadcast_arrays(*rows))


class Affine(Rhs):
    """y' = th0*y + th1.  Covers ozone/tests/test_correctness/test_1_simple/utils.py:6-7
    (th = (-1, 0)) and examples/simple_examples/tutorial.ipynb cell 4 (th = (0.5, -3))."""
    name = 'affine'
    D, P, S = 1, 0, 2

    def f(self, y, p, th):
        return (th[0] * y[0] + th[1])[None]


class QuadDecay(Rhs):
    """ozone/tests/test_correctness/test_3_large/utils.py:23-30:  y' = -(y*(dp + sp))**2,
    one scalar ODE per trajectory, dynamic dp(t) shared, static sp per trajectory."""
    name = 'quad_decay'
    D, P, S = 1, 1, 1

    def f(self, y, p, th):
        return (-(y[0] * (p[0] + th[0])) ** 2)[None]


class General3(Rhs):
    """ozone/tests/test_correctness/test_2_general/utils.py:43-84.
    States packed (y, x, z00, z01, z10, z11); dynamic (a, b, g, e00, e01, e10, e11, coeffs);
    static d.  Profile outputs in insertion order ('z', 'x', 'y') -> 1 + 1 + 4 values,
    field outputs ('y', 'z') -> 1 + 4 values."""
    name = 'general3'
    D, P, S, NPROF, NFIELD = 6, 8, 1, 6, 5

    def f(self, y, p, th):
        yy, xx, z = y[0], y[1], y[2:6]
        a, b, g, e = p[0], p[1], p[2], p[3:7]
        d = th[0]
        z01 = z[1]
        rows = [a * yy - b * yy * xx, g * xx * yy - d * xx]
        for k in range(4):
            rows.append(-z[k] / 3.0 * z01 + yy ** 2 / 2.0 + a * e[k] + xx / 2.0)
        return np.stack(np.broadcast_arrays(*rows))

    def out(self, y, p, th):
        yy, xx, z = y[0], y[1], y[2:6]
        a, e, co = p[0], p[3:7], p[7]
        d = th[0]
        z01, z10 = z[1], z[2]
        prof = [z[3] + 0 * a, xx + (yy / 4.0) * yy + 0 * a]
        for k in range(4):
            prof.append(z[k] * z01 * z10 + e[k] ** 2 + a + d)
        field = [co * yy] + [co * z[k] for k in range(4)]
        return np.stack(np.broadcast_arrays(*prof)), np.stack(np.broadcast_arrays(*field))


class DiffusionReaction(Rhs):
    """ozone/paper_examples/pde_control/pde_control.py:11-81 (ONE trajectory whose state is a grid).
    The cells are the state components (axis 0); zero Dirichlet ghosts (:32-35); the control
    acts where round(0.3 nx) <= j < round(0.7 nx) (:15-16).  Field output J_space uses the
    rebound dx = 1/nx (:62-76): J = (100/nx) sum x^2 + 20 u^2."""
    name = 'diffusion_reaction'
    P, S, NFIELD = 1, 0, 1

    def __init__(self, nx, dx=None, beta_t=16.0, beta_u=2.0, gamma=2.0):
        self.D = self.nx = nx
        self.dx = np.pi / nx if dx is None else dx          # pde_control.py:99
        self.beta_t, self.beta_u, self.gamma = beta_t, beta_u, gamma
        H = np.zeros(nx)
        H[round(0.3 * nx):round(0.7 * nx)] = 1.0
        self.H = H

    def f(self, y, p, th):
        x, u = y, p[0]
        shp = (self.nx,) + (1,) * (x.ndim - 1)
        H = self.H.reshape(shp)
        lap = -2.0 * x
        lap[1:] = lap[1:] + x[:-1]
        lap[:-1] = lap[:-1] + x[1:]
        react = self.beta_t * np.exp(-self.gamma / (1.0 + x)) - self.beta_t * np.exp(-self.gamma)
        return lap / self.dx ** 2 + react + self.beta_u * (H * u - x)

    def out(self, y, p, th):
        u = p[0]
        J = (100.0 / self.nx) * np.sum(y * y, axis=0) + 20.0 * u * u
        shp = np.broadcast_shapes(y.shape[1:], u.shape)
        return np.zeros((0,) + shp), np.broadcast_to(J, shp)[None]

    # analytic VJPs: complex-step columns over 10^6 cells are not an option
    def vjp(self, y, p, th, v):
        x = y
        shp = (self.nx,) + (1,) * (x.ndim - 1)
        H = self.H.reshape(shp)
        lapT = -2.0 * v
        lapT[1:] = lapT[1:] + v[:-1]
        lapT[:-1] = lapT[:-1] + v[1:]
        dreact = self.beta_t * np.exp(-self.gamma / (1.0 + x)) * self.gamma / (1.0 + x) ** 2
        ybar = lapT / self.dx ** 2 + (dreact - self.beta_u) * v
        pbar = np.sum(self.beta_u * H * v, axis=0)[None]
        return ybar, pbar, np.zeros((0,) + v.shape[1:])

    def out_vjp(self, y, p, th, vprof, vfield):
        ybar = (200.0 / self.nx) * y * vfield[0]
        pbar = (40.0 * p[0] * vfield[0])[None]
        return ybar, pbar, np.zeros((0,) + ybar.shape[1:])


REGISTRY = {c.name: c for c in (LotkaVolterra, VanDerPol, LunarAscent, Affine, QuadDecay, General3)}
REGISTRY[DiffusionReaction.name] = DiffusionReaction
