"""Pin the oracle on every known answer the reference's own tests hold for this path
(pins G1-G4).  CPU only."""
import numpy as np
import pytest

from oracle import rhs as R, integrators as I, adjoint as Adj
from oracle.tableaux import tableau, ALL_NAMES
from tests import cases


def _rhs(name, **kw):
    return R.REGISTRY[name](**kw)


def _run(case, method, **kw):
    c = dict(case)
    rhs = _rhs(c.pop('rhs'))
    return rhs, c, I.time_marching(rhs, method, c['y0'], c['h'], c['p'], c['th'], **kw)


def test_G1_tutorial_gl4_trajectory():
    # examples/simple_examples/tutorial.ipynb:161-171 (8 printed digits) + :149 ("converged in 1 iterations")
    gold = [1.0, 0.74364452, 0.47414541, 0.19082879, -0.10701378, -0.42012707, -0.74929402,
            -1.09533772, -1.45912346, -1.8415609, -2.24360632]
    _, _, out = _run(cases.tutorial_inputs(), 'GaussLegendre4')
    assert np.allclose(out['history'][:, 0, 0], gold, rtol=0, atol=5.1e-9)
    assert (out['newton_iters'] == 1).all()


@pytest.mark.parametrize('method', ['RK4', 'Trapezoidal'])
def test_G2_test1_simple(method):
    # ozone/tests/test_correctness/test_1_simple/utils.py:41-52 (same tolerances as the reference)
    rhs, c, out = _run(cases.test1_inputs(), method)
    assert out['history'].sum() == pytest.approx(13.0239779, rel=1e-5)
    g = Adj.tm_adjoint(rhs, method, c['y0'], c['h'], c['p'], c['th'], hist_bar=np.ones_like(out['history']))
    assert np.linalg.norm(g['y0']) == pytest.approx(26.047954954623957, rel=1e-4)
    assert np.linalg.norm(g['h']) == pytest.approx(38.575000915365756, rel=1e-4)


def test_G2_exact_generators():
    # the y0 golden is the exact Trapezoidal value, the h golden the exact RK4 value
    rhs, c, out = _run(cases.test1_inputs(), 'Trapezoidal')
    g = Adj.tm_adjoint(rhs, 'Trapezoidal', c['y0'], c['h'], c['p'], c['th'], hist_bar=np.ones_like(out['history']))
    assert np.linalg.norm(g['y0']) == pytest.approx(26.047954954623957, rel=1e-13)
    rhs, c, out = _run(cases.test1_inputs(), 'RK4')
    g = Adj.tm_adjoint(rhs, 'RK4', c['y0'], c['h'], c['p'], c['th'], hist_bar=np.ones_like(out['history']))
    assert np.linalg.norm(g['h']) == pytest.approx(38.575000915365756, rel=1e-13)


def _test2_totals(res):
    pox, poz, poy = res['profile'][:, 1, 0], res['profile'][:, 0, 0], res['profile'][:, 2:6, 0]
    total = pox[-1] + poz[-1] + res['field'][0, 0] + res['history'][-1, 3, 0] / 2.0
    total2 = np.linalg.norm(poy[-1] + poy[0])
    return total, total2


@pytest.mark.parametrize('method,rel', [('RK4', 1e-13), ('Trapezoidal', 1e-4)])
def test_G3_test2_general_outputs(method, rel):
    # ozone/tests/test_correctness/test_2_general/utils.py:157-162; generated with RK4
    _, _, out = _run(cases.test2_inputs(), method)
    total, total2 = _test2_totals(out)
    assert total == pytest.approx(2.80611596, rel=max(rel, 2e-9))
    assert total2 == pytest.approx(22.27526371713729, rel=rel)


def test_G3_test2_general_jacobian_norms():
    # ozone/tests/test_correctness/test_2_general/utils.py:127-156 -- all 8, at round-off level
    rhs, c, out = _run(cases.test2_inputs(), 'RK4')
    hb = np.zeros_like(out['history']); hb[-1, 3, 0] = 0.5
    pb = np.zeros_like(out['profile']); pb[-1, 1, 0] = 1.0; pb[-1, 0, 0] = 1.0
    fb = np.zeros_like(out['field']); fb[0, 0] = 1.0
    g1 = Adj.tm_adjoint(rhs, 'RK4', c['y0'], c['h'], c['p'], c['th'], history=out['history'],
                        hist_bar=hb, prof_bar=pb, field_bar=fb)
    assert np.linalg.norm(g1['p'][:, 0]) == pytest.approx(0.00574943937426972, rel=1e-10)
    assert np.linalg.norm(g1['y0'][1]) == pytest.approx(1.08155650643438, rel=1e-12)
    assert np.linalg.norm(g1['h']) == pytest.approx(57.652844215047104, rel=1e-12)
    poy = out['profile'][:, 2:6, 0]
    v = poy[-1] + poy[0]
    pb = np.zeros_like(out['profile']); pb[-1, 2:6, 0] = v / np.linalg.norm(v); pb[0, 2:6, 0] = v / np.linalg.norm(v)
    g2 = Adj.tm_adjoint(rhs, 'RK4', c['y0'], c['h'], c['p'], c['th'], history=out['history'], prof_bar=pb)
    assert np.linalg.norm(g2['p'][:, 0]) == pytest.approx(1.754553251836068, rel=1e-12)
    assert np.linalg.norm(g2['y0'][1]) == pytest.approx(0.027324051014, rel=1e-10)
    assert np.linalg.norm(g2['h']) == pytest.approx(33.7649223785, rel=1e-10)
    assert np.linalg.norm(g2['y0'][2:6]) == pytest.approx(9.09442575264, rel=1e-10)
    assert np.linalg.norm(g2['p'][:, 3:7]) == pytest.approx(6.880652370272789, rel=1e-12)


def test_G3_checkpoints_variant():
    # test_2_general/test_solution.py:29-45 runs TimeMarchingCheckpoints(num_checkpoints=5) on the same goldens
    c = cases.test2_inputs()
    rhs = _rhs(c['rhs'])
    out = I.time_marching_checkpoints(rhs, 'RK4', c['y0'], c['h'], c['p'], c['th'], num_checkpoints=5)
    total, total2 = _test2_totals(out)
    assert total2 == pytest.approx(22.27526371713729, rel=1e-13)
    assert out['checkpoints'].shape[0] == 6 and out['num_inner'] == 6
    with pytest.raises(ValueError):      # time_marching_uniform.py:25-26 (29 steps, nc = 5)
        I.checkpoint_layout(29, 5)
    assert I.checkpoint_layout(1024) == (32, 32)


def test_G4_test3_large_exact():
    # ozone/tests/test_correctness/test_3_large/utils.py:55-74 -- exact RK4 fp64 values
    rhs, c, out = _run(cases.test3_inputs(), 'RK4')
    assert out['history'].sum() == pytest.approx(6794.12754072, rel=1e-11)
    g = Adj.tm_adjoint(rhs, 'RK4', c['y0'], c['h'], c['p'], c['th'], history=out['history'],
                       hist_bar=np.ones_like(out['history']))
    assert np.linalg.norm(g['y0']) == pytest.approx(555.6297961976778, rel=1e-14)
    assert np.linalg.norm(g['h']) == pytest.approx(12570.03729508261, rel=1e-14)
    assert np.linalg.norm(g['th']) == pytest.approx(43.805547522344185, rel=1e-14)
    assert np.linalg.norm(g['p']) == pytest.approx(203.55862089571335, rel=1e-14)


def test_G4_test3_trapezoidal_within_reference_tolerance():
    rhs, c, out = _run(cases.test3_inputs(), 'Trapezoidal', tol=1e-7)
    assert out['history'].sum() == pytest.approx(6794.12754072, rel=1e-5)
    g = Adj.tm_adjoint(rhs, 'Trapezoidal', c['y0'], c['h'], c['p'], c['th'], history=out['history'],
                       hist_bar=np.ones_like(out['history']))
    assert np.linalg.norm(g['y0']) == pytest.approx(555.6297961976778, rel=1e-4)
    assert np.linalg.norm(g['h']) == pytest.approx(12570.03729508261, rel=1e-4)
    assert np.linalg.norm(g['th']) == pytest.approx(43.805547522344185, rel=1e-4)
    assert np.linalg.norm(g['p']) == pytest.approx(203.55862089571335, rel=1e-4)


@pytest.mark.parametrize('method', ['RK4', 'Trapezoidal'])
def test_G2_picard_and_collocation_reach_the_same_states(method):
    # test_1_simple/test_solution.py:47-81: the same golden for PicardIteration and Collocation
    c = cases.test1_inputs()
    rhs = _rhs(c['rhs'])
    tm = I.time_marching(rhs, method, c['y0'], c['h'], c['p'], c['th'], tol=1e-13)
    pic = I.picard(rhs, method, c['y0'], c['h'], c['p'], c['th'], tol=1e-13)
    assert pic['converged'].all()
    assert np.allclose(pic['history'], tm['history'], rtol=1e-12, atol=0)
    assert pic['history'].sum() == pytest.approx(13.0239779, rel=1e-5)
    # the time-marching solution is the feasible point of the collocation constraints
    A, b, cc, explicit = tableau(method)
    n = len(c['h'])
    ps = I.interpolate_params(np.zeros((n + 1, 0, 1)), cc)
    Y = np.stack([I.step(rhs, (A, b, cc, explicit), tm['history'][k], c['h'][k], ps[k], c['th'], 1e-14)[1]
                  for k in range(n)])
    c1, c2 = I.collocation_residual(rhs, method, c['y0'], c['h'], Y, tm['history'][1:], c['p'], c['th'])
    assert np.abs(c1).max() < 1e-14 and np.abs(c2).max() < 1e-14


def test_adjoint_matches_complex_step_lotka_volterra():
    # config C1: gradients of sum_t (x + y) w.r.t. a, c, h, x0, y0 (lotka_volterra.py:76-79)
    c = cases.lotka_volterra_inputs(N=1, nt=21)
    rhs = _rhs(c['rhs'])

    def solve(**kw):
        return I.time_marching(rhs, 'RK4', kw['y0'], kw['h'], kw['p'], kw['th'])['profile'].sum()

    out = I.time_marching(rhs, 'RK4', c['y0'], c['h'], c['p'], c['th'])
    g = Adj.tm_adjoint(rhs, 'RK4', c['y0'], c['h'], c['p'], c['th'], history=out['history'],
                       prof_bar=np.ones_like(out['profile']))
    inp = dict(y0=c['y0'], h=c['h'], p=c['p'], th=c['th'])
    for k in ('y0', 'h', 'p', 'th'):
        ref = Adj.complex_step(solve, inp, k)
        assert np.allclose(g[k], ref, rtol=1e-12, atol=1e-14), k


@pytest.mark.parametrize('method', ['GaussLegendre2', 'GaussLegendre4', 'GaussLegendre6', 'RadauII5', 'Lobatto4'])
def test_implicit_adjoint_matches_finite_differences(method):
    c = cases.van_der_pol_inputs(N=3, nt=9)
    c['h'] = np.full(8, 15.0 / 39)          # the config's step size (nt = 40, tf = 15)
    rhs = _rhs(c['rhs'])
    kw = dict(tol=1e-13)
    out = I.time_marching(rhs, method, c['y0'], c['h'], c['p'], c['th'], **kw)
    fb = np.zeros((3, 3)); fb[2] = 1.0
    g = Adj.tm_adjoint(rhs, method, c['y0'], c['h'], c['p'], c['th'], history=out['history'], final_bar=fb, **kw)

    def J(p, h):
        return I.time_marching(rhs, method, c['y0'], h, p, c['th'], **kw)['final'][2].sum()

    rng = np.random.default_rng(1)
    dp, dh = rng.standard_normal(c['p'].shape), rng.standard_normal(c['h'].shape)
    e = 1e-6
    fd = (J(c['p'] + e * dp, c['h'] + e * dh) - J(c['p'] - e * dp, c['h'] - e * dh)) / (2 * e)
    assert (g['p'] * dp).sum() + (g['h'] * dh).sum() == pytest.approx(fd, rel=2e-7)


def test_collocation_vjp_matches_complex_step():
    c = cases.ascent_inputs(N=2, nt=6)
    rhs = _rhs(c['rhs'])
    rng = np.random.default_rng(2)
    n, s, d, N = 5, 4, 7, 2
    Y = rng.standard_normal((n, s, d, N)) * 0.1 + c['y0']
    y = rng.standard_normal((n, d, N)) * 0.1 + c['y0']
    mu1, mu2 = rng.standard_normal((n, s, d, N)), rng.standard_normal((n, d, N))
    g = Adj.colloc_vjp(rhs, 'RK4', c['y0'], c['h'], Y, y, c['p'], c['th'], mu1, mu2)

    def L(**kw):
        # complex-safe re-statement through the residual itself
        c1, c2 = I.collocation_residual(rhs, 'RK4', kw['y0'], kw['h'], kw['Y'], kw['y'], kw['p'], None)
        return (mu1 * c1).sum() + (mu2 * c2).sum()

    e = 1e-6
    for key, base in (('Y', Y), ('y', y), ('p', c['p']), ('h', c['h']), ('y0', c['y0'])):
        dx = rng.standard_normal(base.shape)
        kw = dict(y0=c['y0'], h=c['h'], Y=Y, y=y, p=c['p'])
        kp, km = dict(kw), dict(kw)
        kp[key] = base + e * dx
        km[key] = base - e * dx
        fd = (L(**kp) - L(**km)) / (2 * e)
        assert (g[key] * dx).sum() == pytest.approx(fd, rel=1e-6, abs=1e-9), key


def test_picard_lotka_volterra_converges_and_matches_time_marching():
    # T = 0.9 converges in 15-18 sweeps; result equals the RK4 time-marching solution
    c = cases.lotka_volterra_inputs(N=4, nt=101)
    rhs = _rhs(c['rhs'])
    tm = I.time_marching(rhs, 'RK4', c['y0'], c['h'], c['p'], c['th'])
    pic = I.picard(rhs, 'RK4', c['y0'], c['h'], c['p'], c['th'], tol=1e-12, state_blocks=[slice(0, 1), slice(1, 2)])
    assert pic['converged'].all() and pic['sweeps'].max() < 40
    assert np.allclose(pic['history'], tm['history'], rtol=1e-11, atol=0)
    assert np.allclose(pic['profile'], tm['profile'], rtol=1e-11) and np.allclose(pic['field'], tm['field'], rtol=1e-11)


def test_pde_rhs_matches_reference_formula_and_its_vjp():
    nx = 40
    rhs = R.DiffusionReaction(nx)
    rng = np.random.default_rng(3)
    x = 0.5 + 0.1 * rng.standard_normal((nx, 1))
    u = np.array([[-0.3]])
    # literal re-evaluation of pde_control.py:27-55 for one node
    dx = np.pi / nx
    x0 = np.concatenate([[0.0], x[:-1, 0]]); x1 = np.concatenate([x[1:, 0], [0.0]])
    H = np.zeros(nx); H[round(0.3 * nx):round(0.7 * nx)] = 1
    ref = (x0 - 2 * x[:, 0] + x1) / dx ** 2 + 16.0 * np.exp(1) ** (-2.0 / (1 + x[:, 0])) - 16.0 * np.exp(-2.0) \
        + u[0, 0] * 2.0 * H - 2.0 * x[:, 0]
    assert np.allclose(rhs.f(x, u, np.zeros((0, 1)))[:, 0], ref, rtol=1e-13)
    v = rng.standard_normal((nx, 1))
    Jy = Rhs_cs_jac(rhs, x, u)
    yb, pb, _ = rhs.vjp(x, u, np.zeros((0, 1)), v)
    assert np.allclose(yb[:, 0], Jy.T @ v[:, 0], rtol=1e-12)
    e = 1e-6
    fdp = ((rhs.f(x, u + e, np.zeros((0, 1))) - rhs.f(x, u - e, np.zeros((0, 1)))) / (2 * e) * v).sum()
    assert pb[0, 0] == pytest.approx(fdp, rel=1e-8)


def Rhs_cs_jac(rhs, x, u):
    return R.Rhs.jac(rhs, x, u, np.zeros((0, 1)))[:, :, 0]


def test_all_registered_methods_have_tableaux():
    for name in ALL_NAMES:
        A, b, c, explicit = tableau(name)
        assert A.shape == (len(b), len(b)) and abs(b.sum() - 1.0) < 1e-14
        assert np.allclose(c, A.sum(1))
    assert tableau('RK4')[3] and tableau('RK6')[3] and not tableau('Trapezoidal')[3] and not tableau('Lobatto4')[3]
