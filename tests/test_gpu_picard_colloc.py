"""Parity of the parallel-in-time paths (PicardIteration, Collocation) with the oracle, through the C ABI."""
import numpy as np
import pytest
import torch

from oracle import integrators as I, adjoint as Adj
from ozone_b200 import engine
from tests import cases
from tests.gpu_util import make_spec, dev_inputs, oracle_rhs, relerr, np_, unshare

pytestmark = pytest.mark.gpu


def _picard_pair(case, method, blocks, tol=1e-12, dtype=torch.float64, guess=None):
    spec = make_spec(case, method, dtype, state_blocks=tuple(blocks))
    y0, h, p, th = dev_inputs(case, spec)
    g = None if guess is None else torch.tensor(guess, dtype=dtype, device='cuda')
    out = engine.picard_forward(spec, y0, h, p, th, guess=g, tol=tol, max_sweeps=2000)
    rhs = oracle_rhs(case)
    sl = [slice(a, b) for a, b in zip(blocks[:-1], blocks[1:])]
    ref = I.picard(rhs, method, case['y0'], case['h'], case['p'], case['th'], tol=tol, state_blocks=sl, guess=guess)
    return spec, out, ref


@pytest.mark.parametrize('method', ['RK4', 'ForwardEuler', 'KuttaThirdOrder', 'GaussLegendre4', 'Trapezoidal', 'RK6'])
def test_config4_picard_lotka_volterra_matches_oracle_and_time_marching(method):
    c = cases.lotka_volterra_inputs(N=37, nt=101)
    spec, out, ref = _picard_pair(c, method, (0, 1, 2))
    assert int(out['status'].max()) == 0 and ref['converged'].all()
    assert relerr(np_(out['hist']), ref['history']) < 1e-10
    assert relerr(np_(out['final']), ref['final']) < 1e-10
    assert relerr(np_(out['prof']), ref['profile']) < 1e-10
    assert relerr(np_(out['field']), ref['field']) < 1e-10
    # sweeps-to-converge: same relaxation order as the oracle (block Gauss-Seidel over the states)
    assert np.abs(np_(out['iters']) - ref['sweeps']).max() <= 1
    tm = I.time_marching(oracle_rhs(c), method, c['y0'], c['h'], c['p'], c['th'], tol=1e-13)
    assert relerr(np_(out['hist']), tm['history']) < 1e-10


def test_picard_jacobi_variant_and_user_guess():
    c = cases.lotka_volterra_inputs(N=5, nt=41)
    rng = np.random.default_rng(3)
    guess = 2.0 + 0.1 * rng.standard_normal((40, 2, 5))
    spec, out, ref = _picard_pair(c, 'RK4', (0, 2), guess=guess)
    assert relerr(np_(out['hist']), ref['history']) < 1e-10
    assert np.abs(np_(out['iters']) - ref['sweeps']).max() <= 1


def test_picard_general3_three_state_blocks_with_outputs():
    c = cases.test2_inputs()
    spec, out, ref = _picard_pair(c, 'RK4', (0, 1, 2, 6))
    assert relerr(np_(out['hist']), ref['history']) < 1e-10
    assert relerr(np_(out['prof']), ref['profile']) < 1e-10 and relerr(np_(out['field']), ref['field']) < 1e-10


def test_picard_fp32_and_long_horizon_reports_non_convergence():
    c = cases.lotka_volterra_inputs(N=64, nt=101)
    spec, out, ref = _picard_pair(c, 'RK4', (0, 1, 2), tol=1e-4, dtype=torch.float32)   # fp32 residual floor of 800 unknowns ~ 2e-5
    assert int(out['status'].max()) == 0
    assert relerr(np_(out['hist']), ref['history']) < 2e-5
    # the sweep diverges for long horizons; the kernel must flag it per trajectory, not return NaN silently
    c = cases.lotka_volterra_inputs(N=4, nt=101)
    c['h'] = np.full(100, 0.05)                      # T = 5
    spec = make_spec(c, 'RK4', state_blocks=(0, 1, 2))
    y0, h, p, th = dev_inputs(c, spec)
    out = engine.picard_forward(spec, y0, h, p, th, tol=1e-12, max_sweeps=60)
    assert (np_(out['status']) != 0).all()


def test_picard_gradients_via_time_marching_adjoint():
    c = cases.lotka_volterra_inputs(N=11, nt=31)
    spec = make_spec(c, 'RK4', state_blocks=(0, 1, 2))
    y0, h, p, th = dev_inputs(c, spec, requires_grad=True)
    out = engine.picard(spec, y0, h, p, th, tol=1e-13)
    loss = out['hist'].sum() + 2.0 * out['prof'].sum() + out['field'].sum()
    g = torch.autograd.grad(loss, [y0, h, p, th])
    rhs = oracle_rhs(c)
    ref = I.time_marching(rhs, 'RK4', c['y0'], c['h'], c['p'], c['th'])
    gref = Adj.tm_adjoint(rhs, 'RK4', c['y0'], c['h'], c['p'], c['th'], history=ref['history'],
                          hist_bar=np.ones_like(ref['history']), prof_bar=2 * np.ones_like(ref['profile']),
                          field_bar=np.ones_like(ref['field']))
    for mine, key in zip(g, ('y0', 'h', 'p', 'th')):
        assert relerr(unshare(mine, gref[key]), gref[key]) < 1e-10, key


@pytest.mark.parametrize('method', ['RK4', 'GaussLegendre6', 'ForwardEuler'])
@pytest.mark.parametrize('name', ['lunar_ascent', 'lunar_ascent_tf_per_design', 'van_der_pol', 'lotka_volterra_shared'])
def test_config5_collocation_residual_and_vjp(name, method):
    if name == 'lunar_ascent':
        c = cases.ascent_inputs(N=70, nt=21)
    elif name == 'lunar_ascent_tf_per_design':       # tf ~ U(0.08, 0.12) per design -> per-trajectory steps
        c = cases.ascent_inputs(N=45, nt=15)
        tf = np.random.default_rng(9).uniform(0.08, 0.12, size=45)
        c['h'] = np.tile(tf / 14.0, (14, 1))
    elif name == 'van_der_pol':
        c = cases.van_der_pol_inputs(N=45, nt=12)
    else:
        c = cases.lotka_volterra_inputs(N=33, nt=19)
        c['th'] = np.array([[0.5]])                  # shared static parameter: reduced gradient
    spec = make_spec(c, method)
    y0, h, p, th = dev_inputs(c, spec)
    rhs = oracle_rhs(c)
    n, s, d, N = spec.n, spec.method.num_stages, spec.d, spec.N
    rng = np.random.default_rng(7)
    Y = c['y0'][None, None] + 0.05 * rng.standard_normal((n, s, d, N))
    y = c['y0'][None] + 0.05 * rng.standard_normal((n, d, N))
    mu1, mu2 = rng.standard_normal((n, s, d, N)), rng.standard_normal((n, d, N))
    t = lambda a: torch.tensor(a, dtype=torch.float64, device='cuda')
    c1, c2 = engine.colloc_residual(spec, y0, h, p, th, t(Y), t(y))
    r1, r2 = I.collocation_residual(rhs, method, c['y0'], c['h'], Y, y, c['p'], c['th'])
    assert relerr(np_(c1), r1) < 1e-10 and relerr(np_(c2), r2) < 1e-10
    g = engine.colloc_vjp(spec, y0, h, p, th, t(Y), t(y), t(mu1), t(mu2))
    gref = Adj.colloc_vjp(rhs, method, c['y0'], c['h'], Y, y, c['p'], c['th'], mu1, mu2)
    for k in ('Y', 'y', 'y0', 'h', 'p', 'th'):
        if g[k] is None:
            continue
        assert relerr(unshare(g[k], gref[k]), gref[k]) < 1e-10, k


def test_collocation_feasible_point_is_the_time_marching_solution():
    # test_1_simple/test_solution.py:65-81: the optimiser drives c1 = c2 = 0; the time-marching stages/states satisfy them
    c = cases.van_der_pol_inputs(N=8, nt=10)
    c['h'] = np.full(9, 0.2)
    rhs = oracle_rhs(c)
    from oracle.tableaux import tableau
    method = 'GaussLegendre4'
    tab = tableau(method)
    tm = I.time_marching(rhs, method, c['y0'], c['h'], c['p'], c['th'], tol=1e-14)
    ps = I.interpolate_params(c['p'], tab[2])
    Y = np.stack([I.step(rhs, tab, tm['history'][k], c['h'][k], ps[k], np.zeros((0, 1)), 1e-14)[1] for k in range(9)])
    spec = make_spec(c, method)
    y0, h, p, th = dev_inputs(c, spec)
    t = lambda a: torch.tensor(a, dtype=torch.float64, device='cuda')
    c1, c2 = engine.colloc_residual(spec, y0, h, p, th, t(Y), t(tm['history'][1:]))
    assert float(c1.abs().max()) < 1e-12 and float(c2.abs().max()) < 1e-12
