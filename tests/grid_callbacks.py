"""Grid right-hand sides written as ordinary Python callbacks, for the traced-grid tests.

``pde_control`` is the PDE of the reference's example (ozone/paper_examples/pde_control/pde_control.py:11-81) in this package's
spelling (``tests/test_reference_source_pins.py`` checks it against the reference's own function, traced from its source); ``advection_diffusion`` is a second system that exercises what the first does
not: an asymmetric stencil (upwind advection), a variable diffusivity (per-cell neighbour weights), a per-cell source
weight, a nonlinear reaction without an exponential and a field output with a per-cell weight.  Each comes with a numpy
restatement in the oracle's conventions (oracle.rhs.Rhs, derivatives by complex step)."""
import numpy as np

from ozone_b200 import ops as csdl
from oracle.rhs import Rhs


def pde_control(ozone_vars, dx, nx, zero_bc, pi_bc, beta_t, beta_u, gamma):
    """The PDE of the reference's example (same arguments as its ``ode_function``), written the way a user of this package
    would: x' = (x[k-1] - 2 x[k] + x[k+1]) / dx^2 + beta_t (e^(-gamma / (1 + x)) - e^(-gamma)) + beta_u (H u - x), zero ghost
    cells, H = 1 on the middle 40 % of the grid; J_space = (100 / nx) sum x^2 + 20 u^2."""
    window = np.zeros((nx, 1))
    window[round(0.3 * nx):round(0.7 * nx)] = 1.0
    x = csdl.reshape(ozone_vars.states['x'][0, :], (nx, 1))
    u = ozone_vars.dynamic_parameters['u']
    ghost = np.zeros((1, 1))
    second_difference = (csdl.concatenate((ghost, x[:-1, :])) + (-2.0) * x + csdl.concatenate((x[1:, :], ghost))) / dx ** 2
    reaction = beta_t * (np.exp(1) ** (-gamma * ((1 + x) ** (-1)))) - beta_t * np.exp(-gamma)
    transfer = csdl.expand(csdl.reshape(u[0], (1,)), (nx, 1)) * (beta_u * window) - beta_u * x
    ozone_vars.d_states['x'] = (second_difference + reaction + transfer).reshape(1, -1)
    cells = csdl.reshape(x, (nx,))
    control = csdl.reshape(u[0, :], (1,))
    ozone_vars.field_outputs['J_space'] = ((100.0 / nx) * csdl.sum(cells * cells) + 20.0 * (control * control)).reshape(1, 1)


def advection_fields(nx):
    z = (np.arange(nx) + 0.5) / nx
    return dict(diffusivity=0.6 + 0.4 * np.sin(2 * np.pi * z), source=np.exp(-40.0 * (z - 0.35) ** 2),
                weight=1.0 + z * z)


def advection_diffusion(ozone_vars, nx, dz, speed, rate):
    """c' = D(z) (c[k-1] - 2 c[k] + c[k+1]) / dz^2 - speed (c[k] - c[k-1]) / dz + rate c (1 - c) + s(z) q(t)"""
    fl = advection_fields(nx)
    c = ozone_vars.states['c'][0, :]
    q = ozone_vars.dynamic_parameters['q']
    zero = np.zeros(1)
    left = csdl.concatenate((zero, c[:-1]))
    right = csdl.concatenate((c[1:], zero))
    diffusion = fl['diffusivity'] * (left - 2.0 * c + right) / dz ** 2
    advection = -speed * (c - left) / dz
    reaction = rate * c * (1.0 - c)
    source = csdl.expand(q[0], (nx,)) * fl['source']
    ozone_vars.d_states['c'] = (diffusion + advection + reaction + source).reshape(1, nx)
    ozone_vars.field_outputs['cost'] = (dz * csdl.sum(fl['weight'] * c ** 2) + 0.3 * q[0] * q[0] + 0.1 * q[0]).reshape(1, 1)


class AdvectionDiffusionOracle(Rhs):
    name = 'advection_diffusion'
    P, S, NFIELD = 1, 0, 1

    def __init__(self, nx, dz, speed, rate):
        self.D = self.nx = nx
        self.dz, self.speed, self.rate = dz, speed, rate
        self.fl = advection_fields(nx)

    def _col(self, a, like):
        return a.reshape((self.nx,) + (1,) * (like.ndim - 1))

    def f(self, y, p, th):
        c, q = y, p[0]
        left = np.zeros_like(c)
        left[1:] = c[:-1]
        right = np.zeros_like(c)
        right[:-1] = c[1:]
        D, s = self._col(self.fl['diffusivity'], c), self._col(self.fl['source'], c)
        return (D * (left - 2.0 * c + right) / self.dz ** 2 - self.speed * (c - left) / self.dz
                + self.rate * c * (1.0 - c) + s * q)

    def out(self, y, p, th):
        q = p[0]
        J = self.dz * np.sum(self._col(self.fl['weight'], y) * y * y, axis=0) + 0.3 * q * q + 0.1 * q
        shp = np.broadcast_shapes(y.shape[1:], q.shape)
        return np.zeros((0,) + shp), np.broadcast_to(J, shp)[None]


def dirichlet_heat(ozone_vars, kappa, left, right):
    """T' = kappa (T[k-1] - 2 T[k] + T[k+1]) - 0.5 T^3 with constant boundary values T[-1] = left, T[nx] = right."""
    T = ozone_vars.states['T'][0, :]
    lo = csdl.concatenate((np.full(1, left), T[:-1]))
    hi = csdl.concatenate((T[1:], np.full(1, right)))
    ozone_vars.d_states['T'] = (kappa * (lo - 2.0 * T + hi) - 0.5 * T ** 3).reshape(1, -1)


class DirichletHeatOracle(Rhs):
    name = 'dirichlet_heat'
    P = S = NFIELD = 0

    def __init__(self, nx, kappa, left, right):
        self.D = self.nx = nx
        self.kappa, self.left, self.right = kappa, left, right

    def f(self, y, p, th):
        lo = np.full_like(y, self.left)
        lo[1:] = y[:-1]
        hi = np.full_like(y, self.right)
        hi[:-1] = y[1:]
        return self.kappa * (lo - 2.0 * y + hi) - 0.5 * y ** 3


def two_control_fields(nx):
    z = (np.arange(nx) + 0.5) / nx
    return dict(heater=np.exp(-60.0 * (z - 0.25) ** 2), cooler=np.exp(-60.0 * (z - 0.7) ** 2))


def two_controls(ozone_vars, nx, kappa):
    """x' = kappa (x[k-1] - 2 x[k] + x[k+1]) - x^3 + heat(t) a(z) - cool(t) b(z);  cost = sum x^2 / nx + heat^2 + 0.5 heat cool + 2 cool^2"""
    fl = two_control_fields(nx)
    x = ozone_vars.states['x'][0, :]
    heat, cool = ozone_vars.dynamic_parameters['heat'], ozone_vars.dynamic_parameters['cool']
    zero = np.zeros(1)
    lap = csdl.concatenate((zero, x[:-1])) - 2.0 * x + csdl.concatenate((x[1:], zero))
    ozone_vars.d_states['x'] = (kappa * lap - x ** 3 + csdl.expand(heat[0], (nx,)) * fl['heater']
                                + csdl.expand(cool[0], (nx,)) * (-1.0 * fl['cooler'])).reshape(1, nx)
    ozone_vars.field_outputs['cost'] = (csdl.sum(x * x) / nx + heat[0] * heat[0] + 0.5 * heat[0] * cool[0]
                                        + 2.0 * cool[0] * cool[0]).reshape(1, 1)


class TwoControlsOracle(Rhs):
    name = 'two_controls'
    P, S, NFIELD = 2, 0, 1

    def __init__(self, nx, kappa):
        self.D = self.nx = nx
        self.kappa = kappa
        self.fl = two_control_fields(nx)

    def _col(self, a, like):
        return a.reshape((self.nx,) + (1,) * (like.ndim - 1))

    def f(self, y, p, th):
        lo = np.zeros_like(y)
        lo[1:] = y[:-1]
        hi = np.zeros_like(y)
        hi[:-1] = y[1:]
        return (self.kappa * (lo - 2.0 * y + hi) - y ** 3 + self._col(self.fl['heater'], y) * p[0]
                - self._col(self.fl['cooler'], y) * p[1])

    def out(self, y, p, th):
        J = np.sum(y * y, axis=0) / self.nx + p[0] * p[0] + 0.5 * p[0] * p[1] + 2.0 * p[1] * p[1]
        shp = np.broadcast_shapes(y.shape[1:], p[0].shape)
        return np.zeros((0,) + shp), np.broadcast_to(J, shp)[None]
