"""Parity cases added in round 2: BASELINE sizes that had no test (C5 at 16,384 designs, C3 at n = 1024 / nc = 32),
collocation outputs on the autograd tape, per-trajectory static-parameter gradients under Collocation, the
dtype-aware Newton tolerance, and the fp32 Picard convergence count at 2^20 trajectories."""
import numpy as np
import pytest
import torch

import ozone_b200 as ozone
from oracle import integrators as I, adjoint as Adj
from ozone_b200 import engine, parallel
from ozone_b200.approaches import TimeMarching, Collocation
from ozone_b200.glms import get_integration_method
from ozone_b200.methods import GaussLegendre4, RK4
from tests import cases
from tests.gpu_util import make_spec, dev_inputs, oracle_rhs, relerr, np_, unshare

pytestmark = pytest.mark.gpu
t64 = lambda a: torch.tensor(np.ascontiguousarray(a), dtype=torch.float64, device='cuda')


@pytest.mark.parametrize('method', ['RK4', 'GaussLegendre6'])
def test_config5_full_size_collocation_16384_designs(method):
    N, nt = 16384, 101                                              # BASELINE config C5
    c = cases.ascent_inputs(N=N, nt=nt, seed=11)
    spec = make_spec(c, method)
    y0, h, p, th = dev_inputs(c, spec)
    n, s, d = spec.n, spec.method.num_stages, spec.d
    g = torch.Generator(device='cuda').manual_seed(5)
    rnd = lambda *shape: torch.randn(*shape, dtype=torch.float64, device='cuda', generator=g)
    Y = y0[None, None] + 0.05 * rnd(n, s, d, N)
    y = y0[None] + 0.05 * rnd(n, d, N)
    mu1, mu2 = rnd(n, s, d, N), rnd(n, d, N)
    c1, c2 = engine.colloc_residual(spec, y0, h, p, th, Y, y)
    gr = engine.colloc_vjp(spec, y0, h, p, th, Y, y, mu1, mu2)
    # a 64-design slice (the ragged end of the batch is covered by the small cases) against the oracle
    sl = slice(9000, 9064)
    sub = dict(c, y0=c['y0'][:, sl], p=c['p'][:, :, sl])
    rhs = oracle_rhs(sub)
    Ys, ys, m1, m2 = (np_(a[..., sl]) for a in (Y, y, mu1, mu2))
    r1, r2 = I.collocation_residual(rhs, method, sub['y0'], sub['h'], Ys, ys, sub['p'], None)
    assert relerr(np_(c1[..., sl]), r1) < 1e-10 and relerr(np_(c2[..., sl]), r2) < 1e-10
    gref = Adj.colloc_vjp(rhs, method, sub['y0'], sub['h'], Ys, ys, sub['p'], None, m1, m2)
    for k in ('Y', 'y', 'y0', 'p'):
        assert relerr(np_(gr[k][..., sl]), gref[k]) < 1e-10, k
    # h is shared by the batch: its gradient is a sum over all designs, so compare the slice solved on its own
    ss = make_spec(sub, method)
    gs = engine.colloc_vjp(ss, *dev_inputs(sub, ss), Y[..., sl].contiguous(), y[..., sl].contiguous(),
                           mu1[..., sl].contiguous(), mu2[..., sl].contiguous())
    assert relerr(np_(gs['h']), gref['h']) < 1e-10
    # determinism at full size: the rerun is bit-identical (two-stage reductions, no floating-point atomics)
    c1b, c2b = engine.colloc_residual(spec, y0, h, p, th, Y, y)
    grb = engine.colloc_vjp(spec, y0, h, p, th, Y, y, mu1, mu2)
    assert torch.equal(c1, c1b) and torch.equal(c2, c2b)
    for k in ('Y', 'y', 'y0', 'h', 'p'):
        assert torch.equal(gr[k], grb[k]), k


def test_config3_full_size_grid_1024_steps_32_checkpoints():
    nx, n, nc = 1_000_000, 1024, 32                                 # BASELINE config C3 as written
    ni = n // nc
    m = get_integration_method('RK4')
    cfg = (np.pi / nx, 16.0, 2.0, 2.0, float(round(0.3 * nx)), float(round(0.7 * nx)))
    rng = np.random.default_rng(0)
    dx = np.pi / nx
    y0 = torch.full((nx,), 0.5, dtype=torch.float64, device='cuda')
    h = torch.full((n,), 0.4 * dx * dx, dtype=torch.float64, device='cuda')
    u = t64(rng.uniform(-0.6, 0.0, size=n + 1))
    spec = engine.SolveSpec(rhs_name='diffusion_reaction', method=m, N=1, n=n, h_shared=True, p_shared=True, th_shared=True,
                            num_inner=ni, store_history=False, d_override=nx, rhs_cfg=cfg)
    fwd = engine.tm_forward(spec, y0.reshape(-1, 1), h, u.reshape(-1, 1), None, keep_history=False)
    g = engine.tm_adjoint(spec, y0.reshape(-1, 1), h, u.reshape(-1, 1), None, ckpt=fwd['ckpt'],
                          field_bar=torch.ones(1, 1, dtype=torch.float64, device='cuda'))
    assert torch.isfinite(fwd['final']).all() and torch.isfinite(g['y0']).all() and torch.isfinite(g['p']).all()
    # slab decomposition (4 emulated ranks, halo exchange once per segment) reproduces the owned cells bit for bit
    res = parallel.solve_grid_emulated(m, nx, n, ni, 4, y0, h, u, cfg)
    assert torch.equal(res['final'], fwd['final'][:, 0]) and torch.equal(res['y0_bar'], g['y0'][:, 0])
    assert relerr(np_(res['field']), np_(fwd['field'])) < 1e-12 and relerr(np_(res['u_bar']), np_(g['p'])) < 1e-11
    assert relerr(np_(res['h_bar']), np_(g['h'])) < 1e-11
    # a 4,096-cell window against the oracle over the first checkpoint segment (32 steps).  The window is cut far
    # from the control region's edges; it is run as a grid of its own with halo = 4 stages x 32 steps on either
    # side, so its inner cells see exactly the data the full grid gives them.
    halo, W = 4 * ni, 4096
    lo = 500_000 - W // 2
    ck0, ck1 = fwd['ckpt'][0, :, 0], fwd['ckpt'][1, :, 0]
    assert torch.equal(ck0, y0)
    win = slice(lo - halo, lo + W + halo)
    case = dict(rhs='diffusion_reaction', nx=W + 2 * halo, y0=np_(ck0[win])[:, None], h=np_(h[:ni]), p=np_(u[:ni + 1]).reshape(-1, 1, 1),
                th=None)
    from oracle import rhs as R
    orc = R.DiffusionReaction(W + 2 * halo, dx=np.pi / nx)
    orc.H[:] = 1.0                                                   # the window lies inside the control region (0.3 nx .. 0.7 nx)
    ref = I.time_marching(orc, 'RK4', case['y0'], case['h'], case['p'], None)
    got = np_(ck1[lo:lo + W])
    assert relerr(got, ref['final'][halo:halo + W, 0]) < 1e-12


def test_collocation_outputs_are_on_the_autograd_tape():
    """vector_based.py:33-49: under Collocation profile / field outputs are differentiable expressions of the design
    variables and parameters; checked against the oracle's transposed output Jacobians (complex step)."""
    c = cases.test2_inputs()                                         # general3: 6 profile + 5 field outputs, all nonlinear
    spec = make_spec(c, 'RK4')
    nt, d = spec.nt, spec.d
    rng = np.random.default_rng(2)
    hist_np = c['y0'][None] + 0.1 * rng.standard_normal((nt, d, 1))
    hist = t64(hist_np).requires_grad_(True)
    p = t64(c['p'][:, :, 0]).requires_grad_(True)
    th = t64(c['th'][:, 0]).requires_grad_(True)
    prof, fld = engine.outputs_on_history(spec, hist, p, th)
    rhs = oracle_rhs(c)
    pr, fr = rhs.out(np.moveaxis(hist_np, 0, -1)[:, 0], np.moveaxis(c['p'], 0, -1)[:, 0], c['th'])     # (.., nt)
    assert relerr(np_(prof)[:, :, 0], pr.T) < 1e-13 and relerr(np_(fld)[:, 0], fr.sum(axis=1)) < 1e-13
    wp, wf = rng.standard_normal(prof.shape), rng.standard_normal(fld.shape)
    loss = (prof * t64(wp)).sum() + (fld * t64(wf)).sum()
    gh, gp, gt = torch.autograd.grad(loss, [hist, p, th])
    yb, pb, tb = rhs.out_vjp(np.moveaxis(hist_np, 0, -1)[:, 0], np.moveaxis(c['p'], 0, -1)[:, 0], c['th'],
                             wp[:, :, 0].T, np.repeat(wf, nt, axis=1))
    assert relerr(np_(gh)[:, :, 0], yb.T) < 1e-12
    assert relerr(np_(gp), pb.T) < 1e-12
    assert relerr(np_(gt), tb.sum(axis=1)) < 1e-12


def test_collocation_problem_objective_built_from_field_output_has_gradients():
    # Lotka-Volterra under Collocation through the drop-in API: an objective made of outputs reaches the design variables
    nt = 12
    var = lambda v: torch.tensor(np.asarray(v, dtype=np.float64), device='cuda', requires_grad=True)
    x_0, y_0, cst = var([2.0]), var([2.0]), var([0.5])
    a = var(1.0 + np.arange(nt) / nt / 5.0)
    h = var(np.full(nt - 1, 0.1))
    p = ozone.ODEProblem(RK4(), Collocation())
    p.add_state('y', y_0, store_history=True)
    p.add_state('x', x_0, store_history=True)
    p.add_dynamic_parameter('a', a)
    p.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=h))
    p.set_function(ozone.rhs.lotka_volterra, cst)
    out = p.solve()
    w = torch.arange(1, nt + 1, dtype=torch.float64, device='cuda').reshape(nt, 1)
    obj = (out.profile_outputs['x_plus_y'] * w).sum() + 3.0 * out.field_outputs['field_x'].sum()
    g_state, g_x0, g_y0 = torch.autograd.grad(obj, [p.integrator.state_dv, x_0, y_0])
    # profile = x + y at every grid point (weights w), field = sum_t x (weight 3); packed order (y, x)
    assert torch.allclose(g_state[:, 0, 0], w[1:, 0]) and torch.allclose(g_state[:, 1, 0], w[1:, 0] + 3.0)
    assert float(g_y0) == pytest.approx(1.0) and float(g_x0) == pytest.approx(4.0)


def test_collocation_vjp_per_trajectory_static_parameter_gradient():
    c = cases.lotka_volterra_inputs(N=333, nt=27)                   # c ~ U(0.4, 0.6) per trajectory; 26 steps = 4 time chunks
    method = 'GaussLegendre4'
    spec = make_spec(c, method)
    assert not spec.th_shared
    y0, h, p, th = dev_inputs(c, spec)
    n, s, d, N = spec.n, spec.method.num_stages, spec.d, spec.N
    rng = np.random.default_rng(8)
    Y = c['y0'][None, None] + 0.05 * rng.standard_normal((n, s, d, N))
    y = c['y0'][None] + 0.05 * rng.standard_normal((n, d, N))
    mu1, mu2 = rng.standard_normal((n, s, d, N)), rng.standard_normal((n, d, N))
    g = engine.colloc_vjp(spec, y0, h, p, th, t64(Y), t64(y), t64(mu1), t64(mu2))
    gref = Adj.colloc_vjp(oracle_rhs(c), method, c['y0'], c['h'], Y, y, c['p'], c['th'], mu1, mu2)
    for k in ('Y', 'y', 'y0', 'h', 'p', 'th'):
        assert relerr(unshare(g[k], gref[k]), gref[k]) < 1e-10, k
    # and through autograd
    th_t = th.clone().requires_grad_(True)
    c1, c2 = engine.collocation(spec, y0, h, p, th_t, t64(Y), t64(y))
    gt, = torch.autograd.grad((c1 * t64(mu1)).sum() + (c2 * t64(mu2)).sum(), [th_t])
    assert relerr(np_(gt), gref['th']) < 1e-10


def test_fp32_implicit_problem_with_default_options_converges():
    """the fp64 default tolerance 1e-7 cannot be met by an fp32 residual; the default is dtype-aware."""
    nt, N = 40, 64
    rng = np.random.default_rng(1)
    f32 = lambda v, g=False: torch.tensor(np.asarray(v, dtype=np.float32), device='cuda', requires_grad=g)
    u, h = f32(rng.uniform(-0.75, 1.0, size=(nt, N)), True), f32(np.full(nt - 1, 15.0 / (nt - 1)), True)
    p = ozone.ODEProblem(GaussLegendre4(), TimeMarching())          # no engine options
    p.add_state('x0', f32(np.ones(N)), store_history=True)
    p.add_state('x1', f32(np.ones(N)))
    p.add_state('J', f32(np.zeros(N)), store_final=True)
    p.set_timespan(ozone.timespan.StepVector(start=0.0, step_vector=h))
    p.add_dynamic_parameter('u', u)
    p.set_function(ozone.rhs.van_der_pol)
    out = p.solve()
    assert int(out.status.max()) == 0
    per_step = out.iterations.double().mean().item() / (nt - 1)
    assert 2.0 <= per_step <= 4.0, per_step
    c = cases.van_der_pol_inputs(N=N, nt=nt, seed=1)
    ref = I.time_marching(oracle_rhs(c), 'GaussLegendre4', c['y0'], c['h'], c['p'], None, tol=1e-12)
    assert relerr(np_(out.final_states['J']), ref['final'][2]) < 1e-5
    g, = torch.autograd.grad(out.final_states['J'].sum(), [u])
    assert torch.isfinite(g).all()


def test_config4_fp32_picard_full_batch_convergence_count():
    """2^20 trajectories in fp32 at the bench tolerance: every trajectory must converge.  (Round 1 reported 13 that did
    not: their residual norm sat on the fp32 round-off floor)"""
    N = 1 << 20
    c = cases.lotka_volterra_inputs(N=N, nt=101, seed=0)
    spec = make_spec(c, 'RK4', dtype=torch.float32, state_blocks=(0, 1, 2))
    y0, h, p, th = dev_inputs(c, spec)
    out = engine.picard_forward(spec, y0, h, p, th, tol=1e-4, max_sweeps=2000)
    bad = int((out['status'] != 0).sum())
    assert bad == 0, f'{bad} trajectories not converged; sweeps of those: {out["iters"][out["status"] != 0][:8].tolist()}'
    assert int(out['iters'].max()) < 60
