"""trace_grid.py without a GPU: what a grid callback is turned into, and what is refused (with the reason)."""
import numpy as np
import pytest
import torch

from ozone_b200 import ops
from ozone_b200.trace import diff, leaf
from ozone_b200.trace_grid import TracedGridRHS, leaves_of
from tests import grid_callbacks as G


def _evaluate(e, env):
    """numeric value of an expression DAG; env maps (op, idx) of the leaves to numbers"""
    import math
    memo = {}

    def go(x):
        if id(x) in memo:
            return memo[id(x)]
        if x.op == 'const':
            r = x.val
        elif not x.args:
            r = env[(x.op, x.idx)]
        else:
            a = [go(y) for y in x.args]
            r = {'add': lambda: a[0] + a[1], 'sub': lambda: a[0] - a[1], 'mul': lambda: a[0] * a[1], 'div': lambda: a[0] / a[1],
                 'neg': lambda: -a[0], 'powc': lambda: a[0] ** x.val, 'exp': lambda: math.exp(a[0]), 'sin': lambda: math.sin(a[0]),
                 'cos': lambda: math.cos(a[0]), 'log': lambda: math.log(a[0]), 'sqrt': lambda: math.sqrt(a[0])}[x.op]()
        memo[id(x)] = r
        return r
    return go(e)


def test_the_reference_pde_callback_becomes_a_three_point_model_with_one_coefficient_field():
    nx = 120
    dx = np.pi / nx
    t = TracedGridRHS(G.pde_control, (), dict(dx=dx, nx=nx, zero_bc=0, pi_bc=0, beta_t=16.0, beta_u=2.0, gamma=2.0), 'x', nx, ['u'])
    assert {(l.op, l.idx) for l in leaves_of(t.f)} == {('gx', -1), ('gx', 0), ('gx', 1), ('gc', 0), ('p', 0)}
    tab = t.tracer.coefficient_table()
    H = np.zeros(nx)
    H[round(0.3 * nx):round(0.7 * nx)] = 2.0                       # beta_u * heaviside (pde_control.py:15-16,50)
    assert tab.shape == (1, nx) and np.array_equal(tab[0], H)
    # the recorded expression IS the reference formula (:27-55), cell by cell
    rng = np.random.default_rng(0)
    for _ in range(5):
        xl, x, xr, c, u = rng.uniform(0.1, 1.0, 5)
        env = {('gx', -1): xl, ('gx', 0): x, ('gx', 1): xr, ('gc', 0): c, ('p', 0): u}
        want = (xl - 2 * x + xr) / dx ** 2 + 16.0 * np.exp(-2.0 / (1 + x)) - 16.0 * np.exp(-2.0) + u * c - 2.0 * x
        assert _evaluate(t.f, env) == pytest.approx(want, rel=1e-14)
        assert _evaluate(t.dg, env) == pytest.approx(-2 / dx ** 2 + 16.0 * np.exp(-2.0 / (1 + x)) * 2.0 / (1 + x) ** 2 - 2.0, rel=1e-13)
        assert _evaluate(t.q, env) == pytest.approx(100.0 / nx * x * x, rel=1e-14)       # the rebound dx = 1 / nx (:62-76)
        assert _evaluate(t.r, env) == pytest.approx(20.0 * u * u, rel=1e-14)
    assert t.wr_of_left.is_const() and t.wr_of_left.val == pytest.approx(1 / dx ** 2)
    assert t.spec_info(nx) == dict(d=nx, P=1, S=nx, nprof=0, nfield=1, kind=1)
    src = t.cuda_source()
    assert 'pde_exp(' in src and 'struct GridTraced' in src and 'NCOEF = 1' in src


def test_neighbour_weights_that_vary_from_cell_to_cell_are_shifted_for_the_transposed_stencil():
    nx = 64
    t = TracedGridRHS(G.advection_diffusion, (), dict(nx=nx, dz=1.0 / nx, speed=0.7, rate=1.3), 'c', nx, ['q'])
    fl = G.advection_fields(nx)
    tr = t.tracer
    tab = tr.coefficient_table()
    # (J^T G)_k needs the right-neighbour weight of cell k-1 and the left-neighbour weight of cell k+1: D shifted both ways
    shifts = {sh for base, sh in tr.coefs if np.array_equal(tr.bases[base], fl['diffusivity'])}
    assert shifts == {-1, 0, 1}
    for q, (base, sh) in enumerate(tr.coefs):
        want = np.zeros(nx)
        src = tr.bases[base]
        for k in range(nx):
            if 0 <= k + sh < nx:
                want[k] = src[k + sh]
        assert np.array_equal(tab[q], want)


def _trace(f, nx=50, params=('u',)):
    return TracedGridRHS(f, (), {}, 'x', nx, list(params))


def test_what_the_stencil_kernels_cannot_express_is_refused_with_the_reason():
    z1 = np.zeros(1)

    def five_point(ov):
        x = ov.states['x'][0, :]
        ov.d_states['x'] = (ops.concatenate((np.zeros(2), x[:-2])) + ops.concatenate((x[2:], np.zeros(2))) - 2.0 * x).reshape(1, -1)
    with pytest.raises(NotImplementedError, match='three-point stencil'):
        _trace(five_point)

    def nonlinear_neighbour(ov):
        x = ov.states['x'][0, :]
        left = ops.concatenate((z1, x[:-1]))
        ov.d_states['x'] = (left * left - x).reshape(1, -1)
    with pytest.raises(NotImplementedError, match='neighbour cells must enter linearly'):
        _trace(nonlinear_neighbour)

    def shifted_nonlinear(ov):
        x = ov.states['x'][0, :]
        ov.d_states['x'] = ops.concatenate((z1, ops.exp(x)[:-1])).reshape(1, -1)
    with pytest.raises(NotImplementedError, match='ghost cells'):
        _trace(shifted_nonlinear)

    def wrong_side(ov):
        x = ov.states['x'][0, :]
        ov.d_states['x'] = ops.concatenate((z1, x[1:])).reshape(1, -1)
    with pytest.raises(NotImplementedError, match='ghost cells'):
        _trace(wrong_side)

    def traced_boundary(ov):
        x = ov.states['x'][0, :]
        ov.d_states['x'] = (ops.concatenate((ov.dynamic_parameters['u'][0], x[:-1])) - x).reshape(1, -1)
    with pytest.raises(NotImplementedError, match='only constants or slices'):
        _trace(traced_boundary)

    def single_cell(ov):
        x = ov.states['x'][0, :]
        ov.d_states['x'] = (x * x[3]).reshape(1, -1)
    with pytest.raises(TypeError, match='single cells'):
        _trace(single_cell)

    def parameter_squared(ov):
        x = ov.states['x'][0, :]
        u = ov.dynamic_parameters['u']
        ov.d_states['x'] = (ops.expand(u[0] * u[0], (50,)) - x).reshape(1, -1)
    with pytest.raises(NotImplementedError, match='must enter the state derivative linearly'):
        _trace(parameter_squared)

    def captured_tensor(ov):
        x = ov.states['x'][0, :]
        ov.d_states['x'] = (x * torch.ones(50)).reshape(1, -1)
    with pytest.raises(TypeError, match='tensors'):
        _trace(captured_tensor)

    def with_profile(ov):
        x = ov.states['x'][0, :]
        ov.d_states['x'] = (-x).reshape(1, -1)
        ov.profile_outputs['x2'] = (x * x).reshape(1, -1)
    with pytest.raises(NotImplementedError, match='profile outputs'):
        _trace(with_profile)

    def missing_derivative(ov):
        pass
    with pytest.raises(KeyError):                         # the reference's own check (func_wrapper.py:28-53)
        _trace(missing_derivative)

    with pytest.raises(NotImplementedError, match='at most 4'):
        _trace(lambda ov: None, params=('a', 'b', 'c', 'd', 'e'))


def test_generated_grid_model_compiles_and_registers_without_a_gpu():
    from ozone_b200 import _native as nat
    nx = 80
    t = TracedGridRHS(G.advection_diffusion, (), dict(nx=nx, dz=1.0 / nx, speed=0.7, rate=1.3), 'c', nx, ['q'])
    rid = t.plugin_id(4, False, True, 0)
    assert rid >= 1000
    info = nat.rhs_info(rid)
    assert info['kind'] == 1 and info['P'] == 1 and info['nfield'] == 1 and info['S'] == len(t.tracer.coefs)
    assert t.plugin_id(4, False, True, 0) == rid                       # cached per (stages, dtype)
    with pytest.raises(ValueError, match='explicit methods only'):
        t.plugin_id(2, True, True, 0)


def test_constant_boundary_values_become_a_coefficient_field_at_the_boundary_cells():
    nx = 40

    def dirichlet(ov, left, right):
        x = ov.states['x'][0, :]
        lo = ops.concatenate((np.full(1, left), x[:-1]))
        hi = ops.concatenate((x[1:], np.full(1, right)))
        ov.d_states['x'] = (3.0 * (lo - 2.0 * x + hi)).reshape(1, -1)
    t = TracedGridRHS(dirichlet, (0.7, -1.2), {}, 'x', nx, [])
    tab = t.tracer.coefficient_table()
    want_l, want_r = np.zeros(nx), np.zeros(nx)
    want_l[0], want_r[-1] = 0.7, -1.2
    rows = {tuple(r) for r in tab}
    assert tuple(want_l) in rows and tuple(want_r) in rows
    # cell 0 sees 3 (0.7 - 2 x0 + x1), the last cell 3 (x_{n-2} - 2 x_{n-1} - 1.2), interior cells no constant
    x = np.random.default_rng(3).uniform(-1, 1, nx)
    xe = np.concatenate(([0.0], x, [0.0]))
    for k in (0, 1, nx - 1):
        env = {('gx', -1): xe[k], ('gx', 0): xe[k + 1], ('gx', 1): xe[k + 2]}
        env.update({('gc', q): tab[q, k] for q in range(tab.shape[0])})
        left = 0.7 if k == 0 else x[k - 1]
        right = -1.2 if k == nx - 1 else x[k + 1]
        assert _evaluate(t.f, env) == pytest.approx(3.0 * (left - 2.0 * x[k] + right), rel=1e-14)


def test_reflecting_boundaries_are_an_indicator_field_times_the_boundary_cell_and_periodic_ones_are_refused():
    nx = 30

    def neumann(ov, kappa):
        x = ov.states['x'][0, :]
        lo = ops.concatenate((x[:1], x[:-1]))             # ghost = first cell
        hi = ops.concatenate((x[1:], x[-1:]))             # ghost = last cell
        ov.d_states['x'] = (kappa * (lo - 2.0 * x + hi)).reshape(1, -1)
    t = TracedGridRHS(neumann, (2.5,), {}, 'x', nx, [])
    tab = t.tracer.coefficient_table()
    x = np.random.default_rng(5).uniform(-1, 1, nx)
    xe = np.concatenate(([0.0], x, [0.0]))
    for k in (0, 1, nx - 2, nx - 1):
        env = {('gx', -1): xe[k], ('gx', 0): xe[k + 1], ('gx', 1): xe[k + 2]}
        env.update({('gc', q): tab[q, k] for q in range(tab.shape[0])})
        left = x[0] if k == 0 else x[k - 1]
        right = x[-1] if k == nx - 1 else x[k + 1]
        assert _evaluate(t.f, env) == pytest.approx(2.5 * (left - 2.0 * x[k] + right), rel=1e-13, abs=1e-14)
        # the diagonal the reverse sweep uses sees the mirror term: -2 kappa inside, -kappa at the two ends
        assert _evaluate(t.dg, env) == pytest.approx(-2.5 if k in (0, nx - 1) else -5.0, rel=1e-13)

    def periodic(ov):
        x = ov.states['x'][0, :]
        ov.d_states['x'] = (ops.concatenate((x[-1:], x[:-1])) - x).reshape(1, -1)
    with pytest.raises(NotImplementedError, match='three-point stencil'):
        TracedGridRHS(periodic, (), {}, 'x', nx, [])


def test_several_dynamic_parameters_each_get_their_own_weight_field():
    nx = 48
    t = TracedGridRHS(G.two_controls, (), dict(nx=nx, kappa=3.0), 'x', nx, ['heat', 'cool'])
    assert len(t.du) == 2 and t.spec_info(nx)['P'] == 2
    fl = G.two_control_fields(nx)
    tab = t.tracer.coefficient_table()
    for q, key in enumerate(('heater', 'cooler')):
        l, = leaves_of(t.du[q])
        assert l.op == 'gc' and np.array_equal(tab[l.idx], fl[key] * (1.0 if q == 0 else -1.0))
    src = t.cuda_source()
    assert 'static constexpr int P = 2' in src and 'if (q == 1)' in src and 'u[1]' in src


# ---------------------------------------------------------------- traced expression == eager evaluation of the same callback
def _random_grid_callback(rng, nx):
    """A stencil right-hand side drawn from what trace_grid.py accepts, written with ozone_b200.ops so that the SAME function
    runs eagerly on torch tensors (the ground truth) and on a symbolic grid."""
    cl, cr = rng.uniform(0.5, 2.0, 2)
    fields = [rng.uniform(0.5, 1.5, nx) for _ in range(3)]
    left_kind, right_kind = rng.integers(0, 3, 2)             # 0 zero ghost, 1 constant, 2 reflecting
    lval, rval = rng.uniform(-1.0, 1.0, 2)
    nonlin = int(rng.integers(0, 4))
    var_weights = bool(rng.integers(0, 2))
    two_params = bool(rng.integers(0, 2))

    def f(ov):
        x = ov.states['x'][0, :]
        A = torch.as_tensor if isinstance(x, torch.Tensor) else (lambda a: a)      # eager run: per-cell arrays as tensors
        lpad = {0: np.zeros(1), 1: np.full(1, lval), 2: x[:1]}[int(left_kind)]
        rpad = {0: np.zeros(1), 1: np.full(1, rval), 2: x[-1:]}[int(right_kind)]
        left, right = ops.concatenate((lpad, x[:-1])), ops.concatenate((x[1:], rpad))
        w = A(fields[0]) if var_weights else 1.0
        lin = (cl * left - (cl + cr) * x + cr * right) * w        # tensor operand first: the eager run mixes torch and numpy
        centre = [lambda v: -v ** 3, lambda v: 2.0 * ops.exp(-1.5 / (1.0 + v * v)), lambda v: ops.sin(v) * v,
                  lambda v: v / (2.0 + v * v)][nonlin](x)
        u = ov.dynamic_parameters['u']
        src = ops.expand(u[0], (nx,)) * A(fields[1])
        if two_params:
            src = src + ops.expand(ov.dynamic_parameters['v'][0], (nx,)) * A(0.5 * fields[2])
        ov.d_states['x'] = (lin + centre + src).reshape(1, nx)
        cost = 0.3 * ops.sum(x * x * A(fields[2])) + 0.7 * u[0] * u[0]
        if two_params:
            cost = cost + 0.2 * u[0] * ov.dynamic_parameters['v'][0]
        ov.field_outputs['J'] = cost.reshape(1, 1)
    return f, (['u', 'v'] if two_params else ['u'])


@pytest.mark.parametrize('seed', range(12))
def test_traced_grid_expression_equals_eager_evaluation_of_the_same_callback(seed):
    from ozone_b200.func_wrapper import ODEVars
    rng = np.random.default_rng(100 + seed)
    nx = 24
    f, params = _random_grid_callback(rng, nx)
    t = TracedGridRHS(f, (), {}, 'x', nx, params)
    x = rng.uniform(-0.8, 0.8, nx)
    pv = rng.uniform(-1.0, 1.0, len(params))
    ov = ODEVars({'x': torch.tensor(x).reshape(1, nx)}, {n: torch.tensor([[v]]) for n, v in zip(params, pv)}, 1)
    f(ov)                                                           # eager: plain torch tensors
    want_f = ov.d_states['x'].reshape(-1).numpy()
    want_J = float(ov.field_outputs['J'])
    tab = t.tracer.coefficient_table()
    xe = np.concatenate(([0.0], x, [0.0]))
    got_f, q_sum = np.empty(nx), 0.0
    for k in range(nx):
        env = {('gx', -1): xe[k], ('gx', 0): xe[k + 1], ('gx', 1): xe[k + 2]}
        env.update({('p', q): float(v) for q, v in enumerate(pv)})
        env.update({('gc', q): tab[q, k] for q in range(tab.shape[0])})
        got_f[k] = _evaluate(t.f, env)
        q_sum += _evaluate(t.q, env)
    assert np.allclose(got_f, want_f, rtol=1e-13, atol=1e-13)
    assert q_sum + _evaluate(t.r, {('p', q): float(v) for q, v in enumerate(pv)}) == pytest.approx(want_J, rel=1e-13)
    # what the reverse sweep uses: the transposed stencil row by row against autograd of the eager evaluation
    xt = torch.tensor(x, requires_grad=True)
    ov = ODEVars({'x': xt.reshape(1, nx)}, {n: torch.tensor([[v]]) for n, v in zip(params, pv)}, 1)
    f(ov)
    G = rng.standard_normal(nx)
    want_vjp, = torch.autograd.grad((ov.d_states['x'].reshape(-1) * torch.tensor(G)).sum(), [xt])
    Ge = np.concatenate(([0.0], G, [0.0]))
    for k in range(nx):
        env = {('gx', -1): xe[k], ('gx', 0): xe[k + 1], ('gx', 1): xe[k + 2]}
        env.update({('p', q): float(v) for q, v in enumerate(pv)})
        env.update({('gc', q): tab[q, k] for q in range(tab.shape[0])})
        got = (_evaluate(t.wr_of_left, env) * Ge[k] + _evaluate(t.dg, env) * Ge[k + 1] + _evaluate(t.wl_of_right, env) * Ge[k + 2])
        assert got == pytest.approx(float(want_vjp[k]), rel=1e-11, abs=1e-12)
