"""BASELINE.json's full sizes, checked through properties that do not need the oracle at that size:
batch-composition independence (a trajectory's result does not depend on its neighbours), run-to-run
determinism (no floating-point atomics anywhere), agreement of a sampled sub-batch with the oracle, slab
decomposition of the grid, and the degenerate ends of the size range."""
import numpy as np
import pytest
import torch

from oracle import integrators as I, adjoint as Adj
from ozone_b200 import engine, parallel
from ozone_b200.glms import get_integration_method
from tests import cases
from tests.gpu_util import make_spec, dev_inputs, oracle_rhs, relerr, np_, run_pair

pytestmark = pytest.mark.gpu


def _plan(case, method, grads, tol=1e-7, **kw):
    spec = make_spec(case, method, tol=tol, **kw)
    y0, h, p, th = dev_inputs(case, spec)
    plan = engine.TmPlan(spec, grads=grads, cotangents=('final',))
    plan.y0.copy_(y0); plan.h.copy_(h)
    if p is not None:
        plan.p.copy_(p)
    if th is not None:
        plan.th.copy_(th)
    plan.final_bar.fill_(1.0)
    return plan


def test_config2_full_size_van_der_pol_gl6():
    N = 65536                                                   # BASELINE config C2
    c = cases.van_der_pol_inputs(N=N, nt=40, seed=3)
    plan = _plan(c, 'GaussLegendre6', ('h', 'p'))
    plan.forward(); plan.adjoint(); torch.cuda.synchronize()
    assert int(plan.status.max()) == 0
    final, gp, gh = plan.final.clone(), plan.grad['p'].clone(), plan.grad['h'].clone()
    plan.forward(); plan.adjoint(); torch.cuda.synchronize()    # determinism: bit-identical rerun
    assert torch.equal(final, plan.final) and torch.equal(gp, plan.grad['p']) and torch.equal(gh, plan.grad['h'])
    # a ragged sub-batch (not a multiple of any block size) solved on its own gives the same bits per trajectory
    lo, hi = 12345, 12345 + 1001
    sub = dict(c, y0=c['y0'][:, lo:hi], p=c['p'][:, :, lo:hi])
    ps = _plan(sub, 'GaussLegendre6', ('h', 'p'))
    ps.forward(); ps.adjoint(); torch.cuda.synchronize()
    assert torch.equal(ps.final, final[:, lo:hi]) and torch.equal(ps.grad['p'], gp[:, :, lo:hi])
    # and that sub-batch agrees with the oracle (reference Newton tolerance on both sides)
    rhs = oracle_rhs(sub)
    small = dict(sub, y0=sub['y0'][:, :64], p=sub['p'][:, :, :64])
    ref = I.time_marching(rhs, 'GaussLegendre6', small['y0'], small['h'], small['p'], None)
    assert relerr(np_(ps.final[:, :64]), ref['final']) < 1e-7
    # shared-input gradient = sum over trajectories: the full-batch h_bar equals the sum of 4 quarter-batch ones
    acc = torch.zeros_like(gh)
    for q in range(4):
        part = dict(c, y0=c['y0'][:, q * N // 4:(q + 1) * N // 4], p=c['p'][:, :, q * N // 4:(q + 1) * N // 4])
        pq = _plan(part, 'GaussLegendre6', ('h', 'p'))
        pq.forward(); pq.adjoint()
        acc += pq.grad['h']
    assert relerr(np_(acc), np_(gh)) < 1e-12


def test_config4_full_size_lotka_volterra_rk4_one_million_trajectories():
    N = 1_000_000                                               # BASELINE config C4 (not a power of two)
    c = cases.lotka_volterra_inputs(N=N, nt=101, seed=5)
    plan = _plan(c, 'RK4', ('h', 'p', 'th'))
    plan.forward(); plan.adjoint(); torch.cuda.synchronize()
    assert torch.isfinite(plan.final).all() and torch.isfinite(plan.grad['th']).all()
    sl = slice(N - 77, N)                                       # the ragged tail of the last block
    sub = dict(c, y0=c['y0'][:, sl], th=c['th'][:, sl])
    rhs = oracle_rhs(sub)
    ref = I.time_marching(rhs, 'RK4', sub['y0'], sub['h'], sub['p'], sub['th'])
    assert relerr(np_(plan.hist[:, :, sl]), ref['history']) < 1e-12
    assert relerr(np_(plan.prof[:, :, sl]), ref['profile']) < 1e-12
    g = Adj.tm_adjoint(rhs, 'RK4', sub['y0'], sub['h'], sub['p'], sub['th'], history=ref['history'],
                       final_bar=np.ones_like(ref['final']))
    assert relerr(np_(plan.grad['th'][:, sl]), g['th']) < 1e-10
    gh = plan.grad['h'].clone()
    plan.forward(); plan.adjoint(); torch.cuda.synchronize()
    assert torch.equal(gh, plan.grad['h'])                      # deterministic two-stage reduction over 10^6 trajectories


def test_config3_full_size_grid_slabs():
    nx, n, ni = 1_000_000, 64, 8                                # the full grid, a shorter horizon
    m = get_integration_method('RK4')
    cfg = (np.pi / nx, 16.0, 2.0, 2.0, float(round(0.3 * nx)), float(round(0.7 * nx)))
    rng = np.random.default_rng(0)
    dx = np.pi / nx
    y0 = torch.full((nx,), 0.5, dtype=torch.float64, device='cuda')
    h = torch.full((n,), 0.4 * dx * dx, dtype=torch.float64, device='cuda')
    u = torch.tensor(rng.uniform(-0.6, 0.0, size=n + 1), dtype=torch.float64, device='cuda')
    res = parallel.solve_grid_emulated(m, nx, n, ni, 4, y0, h, u, cfg)
    spec = engine.SolveSpec(rhs_name='diffusion_reaction', method=m, N=1, n=n, h_shared=True, p_shared=True, th_shared=True,
                            num_inner=ni, store_history=False, d_override=nx, rhs_cfg=cfg)
    fwd = engine.tm_forward(spec, y0.reshape(-1, 1), h, u.reshape(-1, 1), None, keep_history=False)
    g = engine.tm_adjoint(spec, y0.reshape(-1, 1), h, u.reshape(-1, 1), None, ckpt=fwd['ckpt'],
                          field_bar=torch.ones(1, 1, dtype=torch.float64, device='cuda'))
    assert torch.equal(res['final'], fwd['final'][:, 0])        # owned cells are bit-identical
    assert torch.equal(res['y0_bar'], g['y0'][:, 0])
    assert relerr(np_(res['field']), np_(fwd['field'])) < 1e-13 and relerr(np_(res['u_bar']), np_(g['p'])) < 1e-12
    # symmetry of the problem about the grid centre survives 64 steps exactly enough (physics sanity at full size)
    f = fwd['final'][:, 0]
    assert float((f - f.flip(0)).abs().max()) < 1e-9


@pytest.mark.parametrize('method', ['RK4', 'GaussLegendre4'])
def test_degenerate_sizes_one_trajectory_one_step(method):
    c = cases.van_der_pol_inputs(N=1, nt=2)
    spec, fwd, ref, g, gref = run_pair(c, method, tol=1e-13)
    assert relerr(np_(fwd['hist']), ref['history']) < 1e-12
    for k in ('y0', 'h', 'p'):
        assert relerr(np_(g[k]).reshape(gref[k].shape), gref[k]) < 1e-10
