"""Parity of the CUDA time-marching path (through the C ABI) with the oracle: states, outputs and
adjoint gradients.  Tolerances are the north star's: 1e-10 relative in fp64, 1e-5 in fp32."""
import numpy as np
import pytest
import torch

from tests import cases
from tests.gpu_util import run_pair, relerr, np_, unshare

pytestmark = pytest.mark.gpu

TOL = {torch.float64: 1e-10, torch.float32: 1e-5}


def _compare(case, method, dtype=torch.float64, tol=1e-7, fwd_tol=None, grad_tol=None, **kw):
    spec, fwd, ref, g, gref = run_pair(case, method, dtype, tol=tol, **kw)
    ft = TOL[dtype] if fwd_tol is None else fwd_tol
    gt = TOL[dtype] if grad_tol is None else grad_tol
    if fwd['hist'] is not None:
        assert relerr(np_(fwd['hist']), ref['history']) < ft
    assert relerr(np_(fwd['final']), ref['final']) < ft
    if fwd['prof'] is not None:
        assert relerr(np_(fwd['prof']), ref['profile']) < ft
    if fwd['field'] is not None:
        assert relerr(np_(fwd['field']), ref['field']) < ft
    assert int(fwd['status'].max()) == 0
    for k in ('y0', 'h', 'p', 'th'):
        if g[k] is None:
            continue
        assert relerr(unshare(g[k], gref[k]), gref[k]) < gt, k
    return spec, fwd, ref, g, gref


@pytest.mark.parametrize('dtype', [torch.float64, torch.float32])
def test_config1_lotka_volterra_rk4_single_trajectory(dtype):
    _compare(cases.lotka_volterra_inputs(N=1), 'RK4', dtype, grad_tol=None if dtype == torch.float64 else 5e-5)


@pytest.mark.parametrize('method', ['ForwardEuler', 'ExplicitMidpoint', 'HeunsMethod', 'RalstonsMethod',
                                    'KuttaThirdOrder', 'RK4', 'RK6'])
def test_every_explicit_method_lotka_volterra_batch(method):
    _compare(cases.lotka_volterra_inputs(N=257, nt=41), method)


def test_lotka_volterra_per_trajectory_parameters_and_steps():
    c = cases.lotka_volterra_inputs(N=100, nt=31, shared_a=False)
    rng = np.random.default_rng(5)
    c['h'] = c['h'][:, None] * (1.0 + 0.1 * rng.random((30, 100)))
    _compare(c, 'RK4')


def test_reference_test3_goldens_on_device():
    # ozone/tests/test_correctness/test_3_large/utils.py:55-74 -- exact RK4 values, now from the GPU
    c = cases.test3_inputs()
    spec, fwd, ref, g, gref = _compare(c, 'RK4')
    assert np_(fwd['hist']).sum() == pytest.approx(6794.12754072, rel=1e-11)
    from ozone_b200 import engine
    from tests.gpu_util import dev_inputs
    y0, h, p, th = dev_inputs(c, spec)
    ones = torch.ones_like(fwd['hist'])
    gg = engine.tm_adjoint(spec, y0, h, p, th, hist=fwd['hist'], hist_bar=ones)
    assert np.linalg.norm(np_(gg['y0'])) == pytest.approx(555.6297961976778, rel=1e-12)
    assert np.linalg.norm(np_(gg['h'])) == pytest.approx(12570.03729508261, rel=1e-12)
    assert np.linalg.norm(np_(gg['th'])) == pytest.approx(43.805547522344185, rel=1e-12)
    assert np.linalg.norm(np_(gg['p'])) == pytest.approx(203.55862089571335, rel=1e-12)


@pytest.mark.parametrize('method', ['RK4', 'Trapezoidal'])
def test_reference_test2_general_three_states_with_outputs(method):
    _compare(cases.test2_inputs(), method, tol=1e-13)


def test_reference_test1_and_tutorial():
    _compare(cases.test1_inputs(), 'RK4')
    _compare(cases.test1_inputs(), 'Trapezoidal', tol=1e-13)
    spec, fwd, *_ = _compare(cases.tutorial_inputs(), 'GaussLegendre4')
    gold = [1.0, 0.74364452, 0.47414541, 0.19082879, -0.10701378, -0.42012707, -0.74929402,
            -1.09533772, -1.45912346, -1.8415609, -2.24360632]     # tutorial.ipynb:161-171
    assert np.allclose(np_(fwd['hist'])[:, 0, 0], gold, rtol=0, atol=5.1e-9)
    assert int(fwd['iters'][0]) == 10                               # "converged in 1 iterations" x 10 steps


@pytest.mark.parametrize('method', ['BackwardEuler', 'ImplicitMidpoint', 'Trapezoidal', 'GaussLegendre2',
                                    'GaussLegendre4', 'GaussLegendre6', 'Lobatto2', 'Lobatto4', 'RadauI3', 'RadauI5',
                                    'RadauII3', 'RadauII5'])
def test_config2_van_der_pol_every_implicit_method(method):
    # tight Newton tolerance on both sides -> agreement at 1e-10
    _compare(cases.van_der_pol_inputs(N=200), method, tol=1e-12)


def test_config2_van_der_pol_reference_newton_tolerance():
    # with the reference's own tolerance (1e-7, time_marching.py:220) both sides stop at the same
    # iteration; quadratic convergence leaves them far closer than 1e-7
    spec, fwd, ref, g, gref = _compare(cases.van_der_pol_inputs(N=200), 'GaussLegendre6', tol=1e-7,
                                       fwd_tol=1e-7, grad_tol=1e-6)
    assert np.array_equal(np_(fwd['iters']), ref['newton_iters'].sum(axis=0))


@pytest.mark.parametrize('method', ['GaussLegendre6', 'Trapezoidal', 'RadauII5'])
def test_config2_adjoint_from_cached_stages_equals_resolve(method):
    # the forward sweep can hand its converged stage values to the reverse sweep (no second Newton solve)
    c = cases.van_der_pol_inputs(N=130)
    _, _, _, g1, gref = _compare(c, method, tol=1e-12, cache_stages=True)
    _, _, _, g2, _ = _compare(c, method, tol=1e-12, cache_stages=False)
    for k in ('y0', 'h', 'p'):
        assert relerr(np_(g1[k]), np_(g2[k])) < 1e-13, k      # same stage values; only FMA contraction may differ


@pytest.mark.parametrize('method,h', [('GaussLegendre2', 2.0), ('BackwardEuler', 1.0)])
def test_newton_matrix_that_needs_a_row_exchange(method, h):
    """With x1 = 0 and h a_11 = 1 the first Newton matrix I - h A (x) J has an exactly zero leading diagonal entry:
    the register elimination must hand over to its pivoting routine (the oracle solves with LAPACK's partial
    pivoting).  Neighbouring trajectories start elsewhere and stay on the exchange-free path.  (Larger stage
    counts would need h >= 4 here, where Newton no longer converges on either side; their 4 x 4 and 6 x 6 systems
    are exercised by the linear problem in test_gpu_traced_callbacks.py.)"""
    N = 37
    y0 = np.zeros((3, N))
    y0[0] = 0.5
    y0[1, 1::2] = np.linspace(0.2, 0.9, len(y0[1, 1::2]))           # every other trajectory: no exchange needed
    c = dict(rhs='van_der_pol', y0=y0, h=np.full(3, h), p=np.full((4, 1, N), 0.1) + np.linspace(0, 0.2, N), th=None)
    spec, fwd, ref, g, gref = run_pair(c, method, tol=1e-13)
    assert int(fwd['status'].max()) == 0
    assert relerr(np_(fwd['hist']), ref['history']) < 1e-10
    for k in ('y0', 'h', 'p'):
        assert relerr(np_(g[k]).reshape(gref[k].shape), gref[k]) < 1e-9, k


def test_config2_fp32():
    _compare(cases.van_der_pol_inputs(N=64), 'GaussLegendre4', torch.float32, tol=1e-5, fwd_tol=2e-5, grad_tol=2e-4)


@pytest.mark.parametrize('method', ['RK4', 'GaussLegendre4'])
def test_lunar_ascent_time_marching(method):
    _compare(cases.ascent_inputs(N=33, nt=21), method, tol=1e-13)


@pytest.mark.parametrize('method,ni', [('RK4', 6), ('RK4', 10), ('GaussLegendre4', 5)])
def test_checkpointed_reverse_sweep_recomputes_the_same_gradients(method, ni):
    # TimeMarchingCheckpoints: only nc+1 boundary states are kept, the reverse sweep recomputes
    # (time_marching_uniform.py:77-125); nc = 5 on the 30-step reference problem (test_2_general/test_solution.py:29-45)
    _compare(cases.test2_inputs(), method, tol=1e-13, num_inner=ni, store_history=False)
    c = cases.lotka_volterra_inputs(N=70, nt=31)
    _compare(c, method, tol=1e-13, num_inner=ni, store_history=False)


def test_zero_step_size_is_legal():
    # test_api_odeproblem.py:57 uses step_vector=[0.0, 1.0]
    c = cases.lotka_volterra_inputs(N=3, nt=3)
    c['h'] = np.array([0.0, 0.1])
    _compare(c, 'RK4')


# ---------------------------------------------------------------------------------------------
# config C3: nonlinear diffusion-reaction grid (pde_control.py), explicit methods, fused stages
@pytest.mark.parametrize('method', ['ForwardEuler', 'HeunsMethod', 'KuttaThirdOrder', 'RK4', 'RK6'])
@pytest.mark.parametrize('nx', [50, 3001])
def test_config3_diffusion_reaction_grid_history_mode(method, nx):
    c = cases.pde_inputs(nx=nx, n=12)
    _compare(c, method)


def test_config3_grid_history_with_a_ragged_last_launch():
    # 13 steps = three launches of four fused steps and one of a single step; 2500 cells = three CTAs
    _compare(cases.pde_inputs(nx=2500, n=13), 'RK4')


@pytest.mark.parametrize('nx,n,ni', [(777, 16, 4), (4100, 24, 6), (50, 9, 3), (2500, 15, 5), (1300, 14, 7)])
def test_config3_checkpointed_grid_adjoint(nx, n, ni):
    # TimeMarchingCheckpoints on the grid: only nc+1 states kept, segments replayed in the reverse sweep
    c = cases.pde_inputs(nx=nx, n=n)
    _compare(c, 'RK4', num_inner=ni, store_history=False)


def test_config3_grid_fp32():
    c = cases.pde_inputs(nx=2000, n=10)
    # states to 1e-5; the step-size gradient sums O(1/dx^2) terms of both signs, fp32 keeps ~3 digits of it
    _compare(c, 'RK4', torch.float32, fwd_tol=1e-5, grad_tol=5e-3)


def test_config3_spatial_slabs_match_the_single_slab_solve():
    # the spatial shard of ozone_b200/parallel.py with 3 ranks emulated on one GPU: owned cells, the field
    # output and every gradient must equal the undivided checkpointed solve (halo 2*s*n_inner, one exchange per segment)
    from ozone_b200 import engine, parallel
    from ozone_b200.glms import get_integration_method
    nx, n, ni = 3000, 12, 3
    c = cases.pde_inputs(nx=nx, n=n)
    cfg = (np.pi / nx, 16.0, 2.0, 2.0, float(round(0.3 * nx)), float(round(0.7 * nx)))
    m = get_integration_method('RK4')
    t = lambda a: torch.tensor(np.ascontiguousarray(a), dtype=torch.float64, device='cuda')
    y0, h, u = t(c['y0'][:, 0]), t(c['h']), t(c['p'][:, 0, 0])
    res = parallel.solve_grid_emulated(m, nx, n, ni, 3, y0, h, u, cfg)
    spec = engine.SolveSpec(rhs_name='diffusion_reaction', method=m, N=1, n=n, h_shared=True, p_shared=True,
                            th_shared=True, num_inner=ni, store_history=False, d_override=nx, rhs_cfg=cfg)
    fwd = engine.tm_forward(spec, y0.reshape(-1, 1), h, u.reshape(-1, 1), None, keep_history=False)
    g = engine.tm_adjoint(spec, y0.reshape(-1, 1), h, u.reshape(-1, 1), None, ckpt=fwd['ckpt'],
                          field_bar=torch.ones(1, 1, dtype=torch.float64, device='cuda'))
    assert relerr(np_(res['final']), np_(fwd['final'][:, 0])) < 1e-13
    assert relerr(np_(res['field']), np_(fwd['field'])) < 1e-13
    assert relerr(np_(res['y0_bar']), np_(g['y0'][:, 0])) < 1e-12
    assert relerr(np_(res['h_bar']), np_(g['h'])) < 1e-12
    assert relerr(np_(res['u_bar']), np_(g['p'])) < 1e-12


def test_pipelined_host_batch_equals_the_single_launch_solve():
    # engine.PipelinedTmSolve: sub-batches on their own streams with pinned host buffers (the bench's e2e leg)
    from ozone_b200 import engine
    from tests.gpu_util import make_spec, dev_inputs
    c = cases.van_der_pol_inputs(N=1000, nt=21)
    spec = make_spec(c, 'GaussLegendre4', tol=1e-12)
    y0, h, p, th = dev_inputs(c, spec)
    plan = engine.TmPlan(spec, grads=('h', 'p'), cotangents=('final',))
    plan.y0.copy_(y0); plan.h.copy_(h); plan.p.copy_(p)
    plan.final_bar.zero_(); plan.final_bar[-1].fill_(1.0)
    plan.forward(); plan.adjoint()
    pipe = engine.PipelinedTmSolve(spec, slices=3, grads=('h', 'p'), cotangents=('final',), upload=('h', 'p'))
    pipe.set_host_inputs(y0=y0.cpu(), h=h.cpu(), p=p.cpu())
    for pl in pipe.plans:
        pl.final_bar.zero_(); pl.final_bar[-1].fill_(1.0)
    for _ in range(2):
        pipe.run()
        torch.cuda.synchronize()
        final, grads = pipe.gather()
        assert torch.equal(final, plan.final.cpu())
        assert torch.equal(grads['p'], plan.grad['p'].cpu())
        assert relerr(np_(grads['h']), np_(plan.grad['h'])) < 1e-13      # same terms, different summation order
    # the same pipeline replayed as a CUDA graph, on NEW host inputs written into the pinned buffers it captured
    pipe.capture()
    p2 = p * 0.5
    plan.p.copy_(p2)
    plan.forward(); plan.adjoint()
    pipe.set_host_inputs(p=p2.cpu())
    for _ in range(2):
        pipe.run()
        torch.cuda.synchronize()
        final, grads = pipe.gather()
        assert torch.equal(final, plan.final.cpu())
        assert torch.equal(grads['p'], plan.grad['p'].cpu())
        assert relerr(np_(grads['h']), np_(plan.grad['h'])) < 1e-13


def test_stream_of_independent_host_batches_keeps_several_solves_in_flight():
    # engine.SolveStream (the bench's e2e leg): three slots with DIFFERENT inputs, used round-robin twice, eagerly and as
    # CUDA graphs; every slot must hold exactly what a plain single-launch solve of its inputs gives
    from ozone_b200 import engine
    from tests.gpu_util import make_spec, dev_inputs
    c = cases.van_der_pol_inputs(N=700, nt=21)
    spec = make_spec(c, 'GaussLegendre4', tol=1e-12)
    y0, h, p, th = dev_inputs(c, spec)
    plan = engine.TmPlan(spec, grads=('h', 'p'), cotangents=('final',))
    plan.y0.copy_(y0); plan.h.copy_(h)
    plan.final_bar.zero_(); plan.final_bar[-1].fill_(1.0)
    q = engine.SolveStream(spec, depth=3, slices=2, grads=('h', 'p'), cotangents=('final',), upload=('h', 'p'))
    assert q.depth == 3
    scale = (1.0, 0.5, -0.25)
    want = []
    for slot, f in enumerate(scale):
        q.set_host_inputs(slot, y0=y0.cpu(), h=h.cpu(), p=(p * f).cpu())
        for pl in q.pipes[slot].plans:
            pl.final_bar.zero_(); pl.final_bar[-1].fill_(1.0)
        plan.p.copy_(p * f)
        plan.forward(); plan.adjoint()
        torch.cuda.synchronize()
        want.append((plan.final.cpu().clone(), plan.grad['p'].cpu().clone(), plan.grad['h'].cpu().clone()))
    for graphs in (False, True):
        if graphs:
            q.capture()
        q.begin()
        slots = [q.submit() for _ in range(6)]
        q.end()
        assert slots == [0, 1, 2, 0, 1, 2]
        torch.cuda.synchronize()
        for slot in range(3):
            final, grads = q.gather(slot)
            assert torch.equal(final, want[slot][0])
            assert torch.equal(grads['p'], want[slot][1])
            assert relerr(np_(grads['h']), np_(want[slot][2])) < 1e-13
