"""The reference's own known-answer tests (ozone/tests/test_correctness/test_{1_simple,2_general,3_large})
re-run through the drop-in API on the GPU: same problem set-up calls, same goldens, same tolerances
(and tighter where the golden is exact).  Derivatives come from torch.autograd through the adjoint
kernels, the analogue of ``sim.compute_totals()`` (ozone/tests/run_ode.py:66-73)."""
import numpy as np
import pytest
import torch

import ozone_b200 as ozone
from ozone_b200.approaches import TimeMarching, TimeMarchingCheckpoints, PicardIteration
from ozone_b200.methods import RK4, Trapezoidal

pytestmark = pytest.mark.gpu
DEV = 'cuda'


def var(value, requires_grad=True):
    return torch.tensor(np.asarray(value, dtype=np.float64), device=DEV, requires_grad=requires_grad)


def jac_norms(outputs, inputs):
    """{(of, wrt): Frobenius norm of d of / d wrt}, as check_derivs computes from compute_totals."""
    res = {}
    for on, o in outputs.items():
        grads = torch.autograd.grad(o.sum(), list(inputs.values()), retain_graph=True, allow_unused=True)
        for (iname, _), g in zip(inputs.items(), grads):
            res[(on, iname)] = 0.0 if g is None else float(torch.linalg.norm(g))
    return res


# ---------------------------------------------------------------- test_1_simple/utils.py:9-52
def build_test1(method, approach, nt=30):
    y_0 = var([0.5])
    h_vec = var(np.ones(nt - 1) * 0.01)
    ode_problem = ozone.ODEProblem(method, approach, newton_tolerance=1e-12)
    ode_problem.add_state('y', initial_condition=y_0, store_history=True)
    ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h_vec))
    ode_problem.set_function(ozone.rhs.affine, -1.0)                  # f: d_states['y'] = -states['y']
    integrated_outputs = ode_problem.solve()
    return {'y_0': y_0, 'h': h_vec}, {'y_out': integrated_outputs.states['y'].sum()}


@pytest.mark.parametrize('approach', [TimeMarching(), PicardIteration()], ids=['TimeMarching', 'Picard'])
@pytest.mark.parametrize('method', [RK4(), Trapezoidal()], ids=['RK4', 'Trapezoidal'])
def test_1_simple(method, approach):
    inputs, outputs = build_test1(method, approach)
    assert float(outputs['y_out']) == pytest.approx(13.0239779, rel=1e-5)                       # utils.py:51-52
    d = jac_norms(outputs, inputs)
    assert d[('y_out', 'y_0')] == pytest.approx(26.047954954623957, rel=1e-4)                  # :44-45
    assert d[('y_out', 'h')] == pytest.approx(38.575000915365756, rel=1e-4)                    # :46-47
    if method.name == 'Trapezoidal':        # the y_0 golden IS the Trapezoidal value, the h golden the RK4 one
        assert d[('y_out', 'y_0')] == pytest.approx(26.047954954623957, rel=1e-11)
    else:
        assert d[('y_out', 'h')] == pytest.approx(38.575000915365756, rel=1e-11)


# ---------------------------------------------------------------- test_2_general/utils.py:6-162
def build_test2(method, approach, nt=31):
    num_times = nt
    y_0, x_0 = var([2.0]), var([2.0])
    z_0 = var([[2.0, 1.0], [-1.0, -3.0]])
    t = np.arange(num_times)
    ai = var(1.0 + t / num_times / 5.0)
    bi = var(0.5 + t / num_times / 5.0)
    gi = var(2.0 + t / num_times / 5.0)
    ei = var(np.array([[0.3, 0.2], [0.1, -2.6]])[None] + (t / num_times / 5.0)[:, None, None])
    di = var([0.5])
    coeffs = var(np.ones(num_times) / num_times)
    h_vec = var(np.ones(num_times - 1) * 0.001)

    ode_problem = ozone.ODEProblem(method, approach, newton_tolerance=1e-12, verbose=False)
    ode_problem.add_state('y', initial_condition=y_0)
    ode_problem.add_state('x', initial_condition=x_0)
    ode_problem.add_state('z', initial_condition=z_0, store_history=True)
    ode_problem.add_dynamic_parameter('a', ai)
    ode_problem.add_dynamic_parameter('b', bi)
    ode_problem.add_dynamic_parameter('g', gi)
    ode_problem.add_dynamic_parameter('e', ei)
    ode_problem.add_dynamic_parameter('coefficients', coeffs)
    ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h_vec))
    ode_problem.set_function(ozone.rhs.general3, di)
    outputs = ode_problem.solve()

    foy = outputs.field_outputs['y']
    pox, poz, poy = outputs.profile_outputs['x'], outputs.profile_outputs['z'], outputs.profile_outputs['y']
    z_int = outputs.states['z']
    assert tuple(z_int.shape) == (num_times, 2, 2) and tuple(poy.shape) == (num_times, 2, 2)
    assert tuple(pox.shape) == (num_times, 1) and tuple(foy.shape) == (1,)
    assert 'y' not in outputs.states and 'x' not in outputs.states          # store_history False
    temp = z_int[-1, 0, 1].reshape(1)
    outs = {'total': pox[-1] + poz[-1] + foy[0] + temp / 2.0,
            'total2': torch.linalg.norm(poy[-1, :, :] + poy[0, :, :])}
    return {'a': ai, 'x_0': x_0, 'h': h_vec, 'z_0': z_0, 'e': ei}, outs


@pytest.mark.parametrize('approach', [TimeMarching(), TimeMarchingCheckpoints(num_checkpoints=5), PicardIteration()],
                         ids=['TimeMarching', 'Checkpoints5', 'Picard'])
@pytest.mark.parametrize('method', [RK4(), Trapezoidal()], ids=['RK4', 'Trapezoidal'])
def test_2_general(method, approach):
    inputs, outputs = build_test2(method, approach)
    assert float(outputs['total']) == pytest.approx(2.80611596, rel=1e-4)                       # utils.py:161
    assert float(outputs['total2']) == pytest.approx(22.27526371713729, rel=1e-4)               # :162
    d = jac_norms(outputs, inputs)
    gold = {('total', 'a'): (0.00574943937426972, 1e-3), ('total', 'x_0'): (1.08155650643438, 1e-4),
            ('total', 'h'): (57.652844215047104, 1e-4), ('total2', 'a'): (1.754553251836068, 1e-3),
            ('total2', 'x_0'): (0.027324051014, 1e-4), ('total2', 'h'): (33.7649223785, 1e-4),
            ('total2', 'z_0'): (9.09442575264, 1e-4), ('total2', 'e'): (6.880652370272789, 1e-4)}   # :127-156
    for key, (val, rel) in gold.items():
        if method.name == 'RK4':             # goldens were generated with RK4: round-off level agreement
            rel = 1e-9
        assert d[key] == pytest.approx(val, rel=rel), key
    if method.name == 'RK4':
        assert float(outputs['total2']) == pytest.approx(22.27526371713729, rel=1e-13)


# ---------------------------------------------------------------- test_3_large/utils.py:6-74
def build_test3(method, approach, nt=30, size=500):
    y_0 = var(0.5 * np.ones(size))
    sp = var(np.linspace(0.1, 1.1, size))
    dp = var(np.linspace(0.5, 0.8, nt).reshape(nt, 1))
    h_vec = var(np.ones(nt - 1) * 0.01)
    ode_problem = ozone.ODEProblem(method, approach, newton_tolerance=1e-12)
    ode_problem.add_state('y', initial_condition=y_0, store_history=True)
    ode_problem.add_dynamic_parameter('dynamic_param', dp)
    ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h_vec))
    ode_problem.set_function(ozone.rhs.quad_decay, sp)            # sp is closure-captured in the reference (:18,30)
    outputs = ode_problem.solve()
    assert tuple(outputs.states['y'].shape) == (nt, size)
    return {'y_0': y_0, 'h': h_vec, 'static_param': sp, 'dynamic_param': dp}, {'y_out': outputs.states['y'].sum()}


@pytest.mark.parametrize('method', [RK4(), Trapezoidal()], ids=['RK4', 'Trapezoidal'])
def test_3_large(method):
    inputs, outputs = build_test3(method, TimeMarching())
    rel_out, rel_d = (1e-11, 1e-12) if method.name == 'RK4' else (1e-5, 1e-4)     # exact RK4 goldens
    assert float(outputs['y_out']) == pytest.approx(6794.12754072, rel=rel_out)                # utils.py:73-74
    d = jac_norms(outputs, inputs)
    assert d[('y_out', 'y_0')] == pytest.approx(555.6297961976778, rel=rel_d)                  # :59-60
    assert d[('y_out', 'h')] == pytest.approx(12570.03729508261, rel=rel_d)                    # :62-63
    assert d[('y_out', 'static_param')] == pytest.approx(43.805547522344185, rel=rel_d)        # :65-66
    assert d[('y_out', 'dynamic_param')] == pytest.approx(203.55862089571335, rel=rel_d)       # :68-69


# ---------------------------------------------------------------- examples/simple_examples/lotka_volterra.py
def test_lotka_volterra_example_script():
    num_times = 10
    x_0, y_0 = var([2.0]), var([2.0])
    a_dynamic = var(1.0 + np.arange(num_times) / num_times / 5.0)
    c_static = var([0.5])
    h = var(np.full(num_times - 1, 0.1))
    ode_problem = ozone.ODEProblem(ozone.methods.RK4(), ozone.approaches.TimeMarching())
    ode_problem.add_state('y', y_0)
    ode_problem.add_state('x', x_0)
    ode_problem.add_dynamic_parameter('a', a_dynamic)
    ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h))
    ode_problem.set_function(ozone.rhs.lotka_volterra, c_static)
    outputs = ode_problem.solve()
    x_plus_y = outputs.profile_outputs['x_plus_y']
    assert tuple(x_plus_y.shape) == (num_times, 1) and not outputs.states and not outputs.final_states
    max_x_plus_y = x_plus_y.max()
    grads = torch.autograd.grad(max_x_plus_y, [a_dynamic, c_static, h, x_0])

    from oracle import rhs as R, integrators as I, adjoint as Adj
    rhs = R.LotkaVolterra()
    p = a_dynamic.detach().cpu().numpy().reshape(num_times, 1, 1)
    ref = I.time_marching(rhs, 'RK4', [[2.0], [2.0]], np.full(9, 0.1), p, [[0.5]])
    assert np.allclose(x_plus_y.detach().cpu().numpy()[:, 0], ref['profile'][:, 0, 0], rtol=1e-12)
    k = int(np.argmax(ref['profile'][:, 0, 0]))
    pb = np.zeros_like(ref['profile']); pb[k] = 1.0
    g = Adj.tm_adjoint(rhs, 'RK4', [[2.0], [2.0]], np.full(9, 0.1), p, [[0.5]], history=ref['history'], prof_bar=pb)
    for mine, theirs in zip(grads, (g['p'][:, 0, 0], g['th'][0], g['h'], g['y0'][1])):
        assert np.allclose(mine.cpu().numpy().ravel(), np.ravel(theirs), rtol=1e-10, atol=1e-13)


# ---------------------------------------------------------------- ozone/paper_examples/pde_control/pde_control.py
def test_pde_control_example_through_the_api_with_checkpoints():
    # build_recorder (pde_control.py:84-143) with nx = 600, nt-1 = 36 steps, TimeMarchingCheckpoints (nc = 6 = int(sqrt(36)))
    nx, n = 600, 36
    dx = np.pi / nx
    dt = 0.4 * dx * dx                                   # the reference refuses dt/dx^2 > 0.5 (:101-103)
    h_vec = var(np.ones(n) * dt)
    x_0 = var(np.ones(nx) * 0.5)
    u = var(np.random.default_rng(0).uniform(-0.6, 0.0, size=n + 1))
    ode_problem = ozone.ODEProblem(ozone.methods.RK4(), TimeMarchingCheckpoints(), verbose=False)
    ode_problem.add_state('x', x_0, store_final=True)
    ode_problem.add_dynamic_parameter('u', u)
    ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h_vec))
    ode_problem.set_function(ozone.rhs.diffusion_reaction, dx=dx, nx=nx, zero_bc=0, pi_bc=0, beta_t=16.0, beta_u=2.0, gamma=2.0)
    outputs = ode_problem.solve()
    assert ode_problem.integrator.num_checkpoints == 6 and ode_problem.integrator.num_inner_steps == 6
    Js = outputs.field_outputs['J_space'] * dt            # the example's objective (:134)
    assert tuple(outputs.final_states['x'].shape) == (nx,) and not outputs.states
    g_u, g_h, g_x0 = torch.autograd.grad(Js.sum(), [u, h_vec, x_0])

    from oracle import rhs as R, integrators as I, adjoint as Adj
    rhs = R.DiffusionReaction(nx)
    p = u.detach().cpu().numpy().reshape(n + 1, 1, 1)
    ref = I.time_marching_checkpoints(rhs, 'RK4', np.full((nx, 1), 0.5), np.full(n, dt), p, None)
    assert float(Js.sum()) == pytest.approx(float(ref['field'][0, 0]) * dt, rel=1e-12)
    g = Adj.tm_adjoint(rhs, 'RK4', np.full((nx, 1), 0.5), np.full(n, dt), p, None, history=ref['history'],
                       field_bar=np.full((1, 1), dt))
    assert np.allclose(g_u.cpu().numpy().ravel(), g['p'].ravel(), rtol=1e-10, atol=1e-14)
    assert np.allclose(g_x0.cpu().numpy().ravel(), g['y0'].ravel(), rtol=1e-10, atol=1e-16)
    # J also depends on h through Js = J_space * dt in the example only via the explicit factor; here dt is a constant
    assert np.allclose(g_h.cpu().numpy().ravel(), g['h'].ravel(), rtol=1e-9, atol=1e-12)


def test_van_der_pol_paper_model_gauss_legendre_batch():
    # van_der_pol/main_model.py:16-64 with a batch of control sequences (config C2 through the public API)
    nt, N, tf = 40, 300, 15.0
    rng = np.random.default_rng(1)
    u = var(rng.uniform(-0.75, 1.0, size=(nt, N)))
    h = var(np.ones(nt - 1) * (tf / (nt - 1)))
    ode_problem = ozone.ODEProblem(ozone.methods.GaussLegendre6(), TimeMarching(), newton_tolerance=1e-12)
    ode_problem.add_state('x0', var(np.ones(N), False), store_history=True)
    ode_problem.add_state('x1', var(np.ones(N), False), store_history=True)
    ode_problem.add_state('J', var(np.zeros(N), False), store_final=True)
    ode_problem.set_timespan(ozone.timespan.StepVector(start=0.0, step_vector=h))
    ode_problem.add_dynamic_parameter('u', u)
    ode_problem.set_function(ozone.rhs.van_der_pol)
    out = ode_problem.solve()
    J_end = out.final_states['J']
    assert tuple(J_end.shape) == (N,) and tuple(out.states['x0'].shape) == (nt, N) and 'J' not in out.states
    assert int(out.status.max()) == 0
    g_u, g_h = torch.autograd.grad(J_end.sum(), [u, h])

    from oracle import rhs as R, integrators as I, adjoint as Adj
    rhs = R.VanDerPol()
    y0 = np.tile(np.array([[1.0], [1.0], [0.0]]), (1, N))
    p = u.detach().cpu().numpy().reshape(nt, 1, N)
    hh = np.full(nt - 1, tf / (nt - 1))
    ref = I.time_marching(rhs, 'GaussLegendre6', y0, hh, p, None, tol=1e-12)
    assert np.allclose(J_end.detach().cpu().numpy(), ref['final'][2], rtol=1e-10)
    fb = np.zeros((3, N)); fb[2] = 1.0
    g = Adj.tm_adjoint(rhs, 'GaussLegendre6', y0, hh, p, None, history=ref['history'], final_bar=fb, tol=1e-12)
    assert np.abs(g_u.cpu().numpy() - g['p'][:, 0]).max() / np.abs(g['p']).max() < 1e-10
    assert np.abs(g_h.cpu().numpy() - g['h']).max() / np.abs(g['h']).max() < 1e-10


def test_picard_jacobi_relaxation_reaches_the_same_fixed_point():
    inputs, outputs = build_test2(RK4(), PicardIteration(relaxation='jacobi'))
    assert float(outputs['total2']) == pytest.approx(22.27526371713729, rel=1e-10)
    with pytest.raises(ValueError):
        PicardIteration(relaxation='sor')


def test_picard_with_user_initial_guess_like_the_van_der_pol_model():
    # van_der_pol/main_model.py:37-39 passes initial_guess=np.linspace(...).reshape(-1, 1) per state (collections.py:125-128)
    nt, tf = 12, 1.5
    u = var(np.full(nt, -0.75))
    h = var(np.ones(nt - 1) * (tf / (nt - 1)))
    sols = []
    for guess in (True, False):
        kw = (lambda a, b: dict(initial_guess=np.linspace(a, b, nt - 1).reshape(-1, 1))) if guess else (lambda a, b: {})
        p = ozone.ODEProblem(RK4(), PicardIteration(tolerance=1e-13))
        p.add_state('x0', var([1.0], False), store_history=True, **kw(1.0, 0.0))
        p.add_state('x1', var([1.0], False), store_history=True, **kw(1.0, 0.0))
        p.add_state('J', var([0.0], False), store_final=True, **kw(0.0, 1.0))
        p.set_timespan(ozone.timespan.StepVector(start=0.0, step_vector=h))
        p.add_dynamic_parameter('u', u)
        p.set_function(ozone.rhs.van_der_pol)
        out = p.solve()
        assert int(out.status.max()) == 0
        sols.append((out.states['x0'].detach(), out.final_states['J'].detach(), int(out.iterations.max())))
    assert torch.allclose(sols[0][0], sols[1][0], rtol=1e-11, atol=0) and torch.allclose(sols[0][1], sols[1][1], rtol=1e-11)
    with pytest.raises(ValueError):                                     # collections.py:92-95: wrong guess shape
        p = ozone.ODEProblem(RK4(), PicardIteration())
        p.add_state('x0', var([1.0], False), initial_guess=np.ones((3, 1)))
        p.set_timespan(ozone.timespan.StepVector(start=0.0, step_vector=h))
        p.set_function(ozone.rhs.affine, -1.0)
        p.solve()
