"""Python right-hand side callbacks, traced and compiled to device functors (ozone_b200/trace.py), on the
reference's own problems: the callbacks below are the reference's (csdl ops spelled with ozone_b200.ops /
array methods) and the answers are the reference's goldens."""
import numpy as np
import pytest
import torch

import ozone_b200 as ozone
from ozone_b200 import ops
from ozone_b200.approaches import TimeMarching, TimeMarchingCheckpoints, PicardIteration, Collocation
from ozone_b200.methods import RK4, Trapezoidal, GaussLegendre6
from tests import reference_callbacks as RC

pytestmark = pytest.mark.gpu


def var(value, requires_grad=True):
    return torch.tensor(np.asarray(value, dtype=np.float64), device='cuda', requires_grad=requires_grad)


def norms(out, inputs):
    g = torch.autograd.grad(out.sum(), list(inputs.values()), retain_graph=True, allow_unused=True)
    return {k: (0.0 if gi is None else float(torch.linalg.norm(gi))) for k, gi in zip(inputs, g)}


@pytest.mark.parametrize('approach', [TimeMarching(), PicardIteration()], ids=['TimeMarching', 'Picard'])
@pytest.mark.parametrize('method', [RK4(), Trapezoidal()], ids=['RK4', 'Trapezoidal'])
def test_1_simple_python_callback(method, approach):
    def f(ozone_vars: ozone.ODEVars):                       # test_1_simple/utils.py:6-7, verbatim
        ozone_vars.d_states['y'] = -ozone_vars.states['y']

    y_0, h_vec = var([0.5]), var(np.ones(29) * 0.01)
    p = ozone.ODEProblem(method, approach, newton_tolerance=1e-12)
    p.add_state('y', initial_condition=y_0, store_history=True)
    p.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h_vec))
    p.set_function(f)
    y_out = p.solve().states['y'].sum()
    assert float(y_out) == pytest.approx(13.0239779, rel=1e-5)
    d = norms(y_out, {'y_0': y_0, 'h': h_vec})
    assert d['y_0'] == pytest.approx(26.047954954623957, rel=1e-4)
    assert d['h'] == pytest.approx(38.575000915365756, rel=1e-4)


def general_f(ozone_vars: ozone.ODEVars, d_arg):
    """test_2_general/utils.py:43-84 with csdl.expand / csdl.reshape spelled ops.expand / ops.reshape."""
    a = ozone_vars.dynamic_parameters['a']
    b = ozone_vars.dynamic_parameters['b']
    g = ozone_vars.dynamic_parameters['g']
    e = ozone_vars.dynamic_parameters['e']
    x, y, z = ozone_vars.states['x'], ozone_vars.states['y'], ozone_vars.states['z']
    n = ozone_vars.num_nodes
    z_adjust1 = ops.expand(ops.reshape(z[:, 0, 1], (n,)), (n, 2, 2), 'i->ijk')
    dy_dt = a * y - b * y * x
    dx_dt = g * x * y - d_arg * x
    dz_dt = z * 0.0
    for i in range(n):
        temp_y = y[i] ** 2
        val = -z[i, :, :] / 3.0 * z_adjust1[i, :, :] + temp_y / 2.0 + a[i] * e[i, :, :] + x[i] / 2.0
        dz_dt = dz_dt.set((i, slice(None), slice(None)), val)
    ozone_vars.d_states['y'] = dy_dt
    ozone_vars.d_states['x'] = dx_dt
    ozone_vars.d_states['z'] = dz_dt
    z_adjust2 = ops.expand(ops.reshape(z[:, 1, 0], (n,)), (n, 2, 2), 'i->ijk')
    ozone_vars.profile_outputs['z'] = ops.reshape(z[:, 1, 1], (n,))
    ozone_vars.profile_outputs['x'] = x + (y / 4.0) * y
    ozone_vars.profile_outputs['y'] = z * z_adjust1 * z_adjust2 + e ** 2 + ops.expand(a.flatten(), (n, 2, 2), 'i->ijk') \
        + ops.expand(d_arg, (n, 2, 2))
    coeffs = ozone_vars.dynamic_parameters['coefficients']
    ozone_vars.field_outputs['y'] = coeffs * y
    ozone_vars.field_outputs['z'] = coeffs[:, 0].expand(z.shape, action='i->ijk') * z


@pytest.mark.parametrize('approach', [TimeMarching(), TimeMarchingCheckpoints(num_checkpoints=5)], ids=['TM', 'Checkpoints5'])
def test_2_general_python_callback(approach):
    nt = 31
    t = np.arange(nt)
    y_0, x_0, z_0 = var([2.0]), var([2.0]), var([[2.0, 1.0], [-1.0, -3.0]])
    ai, bi, gi = var(1.0 + t / nt / 5.0), var(0.5 + t / nt / 5.0), var(2.0 + t / nt / 5.0)
    ei = var(np.array([[0.3, 0.2], [0.1, -2.6]])[None] + (t / nt / 5.0)[:, None, None])
    di, coeffs, h_vec = var([0.5]), var(np.ones(nt) / nt), var(np.ones(nt - 1) * 0.001)
    p = ozone.ODEProblem(RK4(), approach, verbose=False)
    p.add_state('y', initial_condition=y_0)
    p.add_state('x', initial_condition=x_0)
    p.add_state('z', initial_condition=z_0, store_history=True)
    for name, v in (('a', ai), ('b', bi), ('g', gi), ('e', ei), ('coefficients', coeffs)):
        p.add_dynamic_parameter(name, v)
    p.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h_vec))
    p.set_function(general_f, di)
    out = p.solve()
    pox, poz, poy = out.profile_outputs['x'], out.profile_outputs['z'], out.profile_outputs['y']
    total = pox[-1] + poz[-1] + out.field_outputs['y'][0] + out.states['z'][-1, 0, 1].reshape(1) / 2.0
    total2 = torch.linalg.norm(poy[-1] + poy[0])
    assert float(total) == pytest.approx(2.80611596, rel=2e-9)                       # utils.py:161-162 (RK4 goldens)
    assert float(total2) == pytest.approx(22.27526371713729, rel=1e-13)
    inputs = {'a': ai, 'x_0': x_0, 'h': h_vec, 'z_0': z_0, 'e': ei}
    d1, d2 = norms(total, inputs), norms(total2, inputs)
    assert d1['a'] == pytest.approx(0.00574943937426972, rel=1e-9)                   # :127-156
    assert d1['x_0'] == pytest.approx(1.08155650643438, rel=1e-9)
    assert d1['h'] == pytest.approx(57.652844215047104, rel=1e-9)
    assert d2['a'] == pytest.approx(1.754553251836068, rel=1e-9)
    assert d2['x_0'] == pytest.approx(0.027324051014, rel=1e-9)
    assert d2['h'] == pytest.approx(33.7649223785, rel=1e-9)
    assert d2['z_0'] == pytest.approx(9.09442575264, rel=1e-9)
    assert d2['e'] == pytest.approx(6.880652370272789, rel=1e-9)


def test_3_large_python_callback_with_captured_static_parameter():
    # test_3_large/utils.py:6-74: 500 independent ODEs = a batch; `sp` is captured by the closure and differentiated
    nt, size = 30, 500
    y_0 = var(0.5 * np.ones(size))
    sp = var(np.linspace(0.1, 1.1, size))
    dp = var(np.linspace(0.5, 0.8, nt).reshape(nt, 1))
    h_vec = var(np.ones(nt - 1) * 0.01)

    def f(ozone_vars: ozone.ODEVars):
        y = ozone_vars.states['y']
        dpv = ozone_vars.dynamic_parameters['dynamic_param']
        ozone_vars.d_states['y'] = -(y * (dpv + sp)) ** 2

    p = ozone.ODEProblem(RK4(), TimeMarching(), batched=True)
    p.add_state('y', initial_condition=y_0, store_history=True)
    p.add_dynamic_parameter('dynamic_param', dp)
    p.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h_vec))
    p.set_function(f)
    y_out = p.solve().states['y'].sum()
    assert float(y_out) == pytest.approx(6794.12754072, rel=1e-11)                   # utils.py:73-74, exact RK4 values
    d = norms(y_out, {'y_0': y_0, 'h': h_vec, 'static_param': sp, 'dynamic_param': dp})
    assert d['y_0'] == pytest.approx(555.6297961976778, rel=1e-12)
    assert d['h'] == pytest.approx(12570.03729508261, rel=1e-12)
    assert d['static_param'] == pytest.approx(43.805547522344185, rel=1e-12)
    assert d['dynamic_param'] == pytest.approx(203.55862089571335, rel=1e-12)


def test_van_der_pol_python_callback_equals_the_registered_functor():
    # van_der_pol/main_model.py:6-14 (tests/reference_callbacks.py, pinned on the reference's file by
    # tests/test_reference_source_pins.py); implicit GL6 -> Newton in registers, J detected as a quadrature state
    ode_func = RC.van_der_pol

    nt, N = 40, 256
    rng = np.random.default_rng(4)
    uval = rng.uniform(-0.75, 1.0, size=(nt, N))
    res = []
    for func, opts in ((ode_func, dict(batched=True)), (ozone.rhs.van_der_pol, {})):
        u, h = var(uval), var(np.ones(nt - 1) * (15.0 / (nt - 1)))
        p = ozone.ODEProblem(GaussLegendre6(), TimeMarching(), newton_tolerance=1e-12, **opts)
        p.add_state('x0', var(np.ones(N), False), store_history=True)
        p.add_state('x1', var(np.ones(N), False))
        p.add_state('J', var(np.zeros(N), False), store_final=True)
        p.set_timespan(ozone.timespan.StepVector(start=0.0, step_vector=h))
        p.add_dynamic_parameter('u', u)
        p.set_function(func)
        out = p.solve()
        J = out.final_states['J']
        g = torch.autograd.grad(J.sum(), [u, h])
        res.append((J.detach(), out.states['x0'].detach(), g[0], g[1], out.iterations))
    if isinstance(ozone.rhs.van_der_pol, object):
        assert p.integrator.spec.traced is None
    for a, b in zip(res[0][:4], res[1][:4]):
        assert float((a - b).abs().max() / b.abs().max()) < 1e-12
    assert torch.equal(res[0][4], res[1][4])                                         # same Newton iteration counts


@pytest.mark.parametrize('approach', ['TimeMarching', 'Collocation'])
def test_lunar_ascent_python_callback_equals_the_registered_functor(approach):
    # ascent_system/main_model.py:6-21 traced vs the hand-written RhsLunarAscent: states, constraints and gradients
    from tests import cases
    nt, N = 21, 70
    case = cases.ascent_inputs(N=N, nt=nt, seed=3)
    names = ('rx', 'ry', 'rz', 'Vx', 'Vy', 'Vz', 'm')
    res = []
    for func, opts in ((RC.lunar_ascent, dict(batched=True)), (ozone.rhs.lunar_ascent, {})):
        u = [var(case['p'][:, i, :]) for i in range(3)]
        tf = var([0.1])
        h = (tf / (nt - 1)).expand(nt - 1)
        p = ozone.ODEProblem(RK4(), TimeMarching() if approach == 'TimeMarching' else Collocation(), **opts)
        for i, nm in enumerate(names):
            p.add_state(nm, var(case['y0'][i], False), store_history=True, store_final=True)
        for nm, ui in zip(('ux', 'uy', 'uz'), u):
            p.add_dynamic_parameter(nm, ui)
        p.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h))
        p.set_function(func)
        out = p.solve()
        if approach == 'TimeMarching':
            obj = sum((out.final_states[nm] ** 2).sum() for nm in names)
            vals = [out.states[nm].detach() for nm in names]
        else:
            integ = p.integrator
            obj = (integ.c1 ** 2).sum() + (integ.c2 * torch.arange(1, integ.c2.numel() + 1, device='cuda',
                                                                    dtype=torch.float64).reshape(integ.c2.shape)).sum()
            vals = [integ.c1.detach(), integ.c2.detach()]
        res.append((vals, torch.autograd.grad(obj, u + [tf])))
    for a, b in zip(res[0][0], res[1][0]):
        assert float((a - b).abs().max()) <= 1e-13 * max(1.0, float(b.abs().max()))
    for a, b in zip(res[0][1], res[1][1]):
        assert float((a - b).abs().max()) <= 1e-13 * max(1.0, float(b.abs().max()))


def test_lotka_volterra_example_python_callback_and_collocation_constraints():
    ode_function = RC.lotka_volterra              # examples/simple_examples/lotka_volterra.py:51-62

    num_times = 10
    outs = {}
    for name, func in (('py', ode_function), ('dev', ozone.rhs.lotka_volterra)):
        x_0, y_0, c = var([2.0]), var([2.0]), var([0.5])
        a = var(1.0 + np.arange(num_times) / num_times / 5.0)
        h = var(np.full(num_times - 1, 0.1))
        p = ozone.ODEProblem(RK4(), TimeMarching())
        p.add_state('y', y_0)
        p.add_state('x', x_0)
        p.add_dynamic_parameter('a', a)
        p.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h))
        p.set_function(func, c)
        o = p.solve()
        m = o.profile_outputs['x_plus_y'].max() + o.field_outputs['field_x'].sum()
        outs[name] = (o.profile_outputs['x_plus_y'].detach(), torch.autograd.grad(m, [a, c, h, x_0]))
    assert torch.allclose(outs['py'][0], outs['dev'][0], rtol=1e-13, atol=0)
    for g1, g2 in zip(outs['py'][1], outs['dev'][1]):
        assert torch.allclose(g1, g2, rtol=1e-11, atol=1e-14)

    # the same callback under Collocation: constraints vanish at the time-marching solution of a traced RHS too
    x_0, y_0, c = var([2.0]), var([2.0]), var([0.5])
    a = var(1.0 + np.arange(num_times) / num_times / 5.0)
    h = var(np.full(num_times - 1, 0.1))
    p = ozone.ODEProblem(RK4(), Collocation())
    p.add_state('y', y_0, store_history=True)
    p.add_state('x', x_0, store_history=True)
    p.add_dynamic_parameter('a', a)
    p.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h))
    p.set_function(ode_function, c)
    o = p.solve()
    integ = p.integrator
    assert set(integ.constraints) == {'y_stage_c', 'y_state_c', 'x_stage_c', 'x_state_c'}   # collocation.py:53-54
    assert tuple(o.states['x'].shape) == (num_times, 1)
    gY, = torch.autograd.grad((integ.c1 ** 2).sum() + (integ.c2 ** 2).sum(), [integ.stage_dv])
    assert torch.isfinite(gY).all() and float(gY.abs().max()) > 0


def test_repeated_solves_reuse_one_plugin_registration():
    from ozone_b200 import trace

    def f(v):
        v.d_states['y'] = -2.0 * v.states['y']
    ids = set()
    for _ in range(5):
        p = ozone.ODEProblem(RK4(), TimeMarching())
        p.add_state('y', var([1.0]), store_final=True)
        p.set_timespan(ozone.timespans.StepVector(0.0, var(np.full(10, 0.05))))
        p.set_function(f)
        out = p.solve()
        ids.add(p.integrator.spec.traced.plugin_id(4, False, True, 0))
        assert float(out.final_states['y']) == pytest.approx(np.exp(-1.0), rel=1e-6)
    assert len(ids) == 1


@pytest.mark.parametrize('method_name', ['BackwardEuler', 'GaussLegendre4', 'GaussLegendre6'])
def test_linear_system_whose_newton_matrix_needs_row_exchanges(method_name):
    """y' = J y with J_00 = 1 / (h a_11): the leading entry of the Newton matrix I - h A (x) J vanishes while the
    entries under it are O(1), so the register elimination has to leave its exchange-free path.  The answer is
    the stability function applied n times, computed here with LAPACK (np.linalg.solve)."""
    from ozone_b200 import methods
    from ozone_b200.glms import get_integration_method
    m = get_integration_method(method_name)
    A, B = np.asarray(m.A, dtype=np.float64), np.asarray(m.B, dtype=np.float64).reshape(-1)
    s, h, n = A.shape[0], 0.1, 3
    J = np.array([[1.0 / (h * A[0, 0]), 1.0], [1.0, 0.0]])

    def f(v: ozone.ODEVars):
        a, b = v.states['a'], v.states['b']
        v.d_states['a'] = J[0, 0] * a + J[0, 1] * b
        v.d_states['b'] = J[1, 0] * a + J[1, 1] * b

    a_0, b_0, h_vec = var([1.0]), var([0.5]), var(np.full(n, h))
    p = ozone.ODEProblem(getattr(methods, method_name)(), TimeMarching())
    p.add_state('a', initial_condition=a_0, store_final=True)
    p.add_state('b', initial_condition=b_0, store_final=True)
    p.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h_vec))
    p.set_function(f)
    out = p.solve()
    total = out.final_states['a'].sum() + 2.0 * out.final_states['b'].sum()
    # one step: (I - h A (x) J) Y = 1 (x) y ;  y+ = y + h (B (x) J) Y  =>  y+ = R y
    M = np.eye(2 * s) - h * np.kron(A, J)
    R = np.eye(2) + h * np.kron(B[None, :], J) @ np.linalg.solve(M, np.kron(np.ones((s, 1)), np.eye(2)))
    assert abs(M[0, 0]) < 1e-12 < 0.1 * np.abs(M[1:, 0]).max()        # the situation this test is about
    Rn = np.linalg.matrix_power(R, n)
    yn = Rn @ np.array([1.0, 0.5])
    assert float(out.final_states['a']) == pytest.approx(yn[0], rel=1e-10)
    assert float(out.final_states['b']) == pytest.approx(yn[1], rel=1e-10)
    ga, gb = torch.autograd.grad(total, [a_0, b_0])
    gref = Rn.T @ np.array([1.0, 2.0])
    assert float(ga) == pytest.approx(gref[0], rel=1e-9) and float(gb) == pytest.approx(gref[1], rel=1e-9)
