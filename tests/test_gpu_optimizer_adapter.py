"""{f, c, df, dc} for an optimiser on BASELINE config C5's model: the lunar-ascent trajectory optimisation exactly as
the reference builds it (ozone/paper_examples/ascent_system/main_model.py:51-125: controls ux, uy, uz and the final
time tf are design variables, tf is the objective, three orbital conditions on the final states, |u|^2 = 1 at every
time point, a lower bound on the final mass), for a batch of independent designs under Collocation."""
import numpy as np
import pytest
import torch

import ozone_b200 as ozone
from oracle import adjoint as Adj
from ozone_b200 import engine
from ozone_b200.approaches import Collocation
from ozone_b200.methods import RK4
from ozone_b200.optimize import CollocationProblem
from tests import cases
from tests.gpu_util import oracle_rhs, np_, relerr

pytestmark = pytest.mark.gpu
NT, M_SCALE = 21, 15103.
NAMES = ('rx', 'ry', 'rz', 'Vx', 'Vy', 'Vz', 'm')


def orbital_conditions(fs, a=1.04603, ecc=0.0385, inc=np.pi / 2):            # main_model.py:23-49, batched over the last axis
    r = torch.stack([fs['rx'].reshape(-1), fs['ry'].reshape(-1), fs['rz'].reshape(-1)])      # (3, N)
    V = torch.stack([fs['Vx'].reshape(-1), fs['Vy'].reshape(-1), fs['Vz'].reshape(-1)])
    rxV = torch.linalg.cross(r, V, dim=0)
    hh = np.sqrt(a * (1 - ecc ** 2))
    c1 = (rxV * rxV).sum(0) - a * (1 - ecc ** 2)
    c2 = (V * V).sum(0) / 2. - 1. / torch.linalg.norm(r, dim=0) + 1. / (2 * a)
    c3 = rxV[2] - hh * np.cos(inc)
    return c1, c2, c3


def ascent_model(dv, collocation_design_values=None):
    N = dv['ux'].shape[-1]
    ic = torch.tensor(cases.ASCENT_IC, dtype=torch.float64, device='cuda')
    tf = dv['tf']                                                             # (N,)
    h = (tf / (NT - 1)).unsqueeze(0).expand(NT - 1, N)                        # one step vector per design
    p = ozone.ODEProblem(RK4(), Collocation(), collocation_design_values=collocation_design_values)
    for i, nm in enumerate(NAMES):
        p.add_state(nm, ic[i].expand(N).clone(), store_final=True)             # scalar state, N designs
    for nm in ('ux', 'uy', 'uz'):
        p.add_dynamic_parameter(nm, dv[nm])                                   # (NT, N)
    p.set_timespan(ozone.timespan.StepVector(start=0.0, step_vector=h))
    p.set_function(ozone.rhs.lunar_ascent)
    out = p.solve()
    c1, c2, c3 = orbital_conditions(out.final_states)
    cons = {'c1': c1.reshape(1, N), 'c2': c2.reshape(1, N), 'c3': c3.reshape(1, N),
            'cu': dv['ux'] ** 2 + dv['uy'] ** 2 + dv['uz'] ** 2 - 1.,       # (NT, N)
            'm_final': out.final_states['m'].reshape(1, N)}
    return p, tf, cons


def _setup(N=64, seed=0):
    rng = np.random.default_rng(seed)
    t = lambda a: torch.tensor(a, dtype=torch.float64, device='cuda')
    dvs = {'ux': t(rng.uniform(-1, 1, (NT, N))), 'uy': t(rng.uniform(-1, 1, (NT, N))), 'uz': t(rng.uniform(-1, 1, (NT, N))),
           'tf': t(rng.uniform(0.08, 0.12, N))}
    return CollocationProblem(ascent_model, dvs), dvs, rng


def test_ascent_optimisation_problem_f_c_df_dc_for_64_designs():
    nlp, dvs, rng = _setup()
    r0 = nlp.compute_optimization_derivatives()
    integ = r0['problem'].integrator
    n, s, d, N = integ.spec.n, 4, 7, 64
    # perturbed collocation unknowns (the initial guess is the broadcast initial condition)
    Y = integ.stage_dv.detach() + 0.02 * torch.randn(n, s, d, N, dtype=torch.float64, device='cuda')
    y = integ.state_dv.detach() + 0.02 * torch.randn(n, d, N, dtype=torch.float64, device='cuda')
    r = nlp.compute_optimization_derivatives(stage_dv=Y, state_dv=y)
    assert torch.equal(r['f'], dvs['tf'])                                     # objective = final time
    assert torch.equal(r['df']['tf'], torch.ones(N, dtype=torch.float64, device='cuda'))
    assert float(r['df']['ux'].abs().max()) == 0.0 and float(r['df']['stage_dv'].abs().max()) == 0.0
    assert set(r['c']) == {'c1', 'c2', 'c3', 'cu', 'm_final', 'stage_c', 'state_c'}
    # scalers as the reference computes them (collections.py:204-209): 1 / mean|guess| per state
    for i, nm in enumerate(NAMES):
        assert r['scalers'][nm]['stage'] == pytest.approx(1.0 / abs(cases.ASCENT_IC[i]), rel=1e-12)

    # ---- collocation block Jacobian of two designs against the oracle's VJP and against finite differences
    blocks = r['dc']['collocation']
    spec = integ.spec
    y0, h, p = integ.y0.detach(), integ.h.detach(), integ.p.detach()
    for i in (0, 37):
        J, cols = blocks.dense(i)
        sub = dict(rhs='lunar_ascent', y0=np_(y0[:, i:i + 1]), h=np_(h[:, i]), p=np_(p[:, :, i:i + 1]), th=None)
        mu1, mu2 = rng.standard_normal((n, s, d, 1)), rng.standard_normal((n, d, 1))
        g = Adj.colloc_vjp(oracle_rhs(sub), 'RK4', sub['y0'], sub['h'], np_(Y[..., i:i + 1]), np_(y[..., i:i + 1]), sub['p'], None,
                           mu1, mu2)
        mu = np.concatenate([mu1.reshape(-1), mu2.reshape(-1)])
        JT = J.T @ mu
        sl = lambda k: slice(cols[k][0], cols[k][0] + cols[k][1])
        assert relerr(JT[sl('Y')], g['Y'].reshape(-1)) < 1e-10 and relerr(JT[sl('y')], g['y'].reshape(-1)) < 1e-10
        assert relerr(JT[sl('y0')], g['y0'].reshape(-1)) < 1e-10 and relerr(JT[sl('h')], g['h'].reshape(-1)) < 1e-10
        assert relerr(JT[sl('p')], g['p'].reshape(-1)) < 1e-10
    # the same transposed product for all 64 designs at once is the VJP kernel
    mu1, mu2 = torch.randn_like(Y), torch.randn_like(y)
    gk = engine.colloc_vjp(spec, y0, h, p, None, Y, y, mu1, mu2)
    J, cols = blocks.dense(5)
    mu = np.concatenate([np_(mu1[..., 5]).reshape(-1), np_(mu2[..., 5]).reshape(-1)])
    assert relerr((J.T @ mu)[cols['Y'][0]:cols['Y'][0] + cols['Y'][1]], np_(gk['Y'][..., 5]).reshape(-1)) < 1e-10

    # ---- user constraints: rows against central differences of the model (design 3, a few directions)
    i, e = 3, 1e-6
    dcu = r['dc']['user']
    assert tuple(dcu['cu']['ux'].shape) == (NT, NT, N)                        # row k depends on ux[k] only
    k = 7
    assert float(dcu['cu']['ux'][k, k, i]) == pytest.approx(2.0 * float(dvs['ux'][k, i]), rel=1e-12)
    assert float(dcu['cu']['ux'][k, k - 1, i]) == 0.0
    # orbital conditions depend on the final state only: d c / d state_dv is nonzero in the last step only
    gs = dcu['c2']['state_dv'][0]                                             # (n, d, N)
    assert float(gs[:-1].abs().max()) == 0.0 and float(gs[-1].abs().max()) > 0.0
    yp, ym = y.clone(), y.clone()
    dirn = torch.randn(d, dtype=torch.float64, device='cuda')
    yp[-1, :, i] += e * dirn
    ym[-1, :, i] -= e * dirn
    cp = nlp.compute_optimization_derivatives(stage_dv=Y, state_dv=yp)['c']
    cm = nlp.compute_optimization_derivatives(stage_dv=Y, state_dv=ym)['c']
    for cname in ('c1', 'c2', 'c3', 'm_final'):
        fd = float((cp[cname][0, i] - cm[cname][0, i]) / (2 * e))
        an = float((dcu[cname]['state_dv'][0][-1, :, i] * dirn).sum())
        assert an == pytest.approx(fd, rel=1e-6, abs=1e-9), cname
    # tf enters the collocation constraints through h = tf / (nt - 1): d c / d tf = sum_k (d c / d h_k) / (nt - 1)
    tfp, tfm = dvs['tf'].clone(), dvs['tf'].clone()
    tfp[i] += e
    tfm[i] -= e
    cp = nlp.compute_optimization_derivatives(values={'tf': tfp}, stage_dv=Y, state_dv=y)['c']
    cm = nlp.compute_optimization_derivatives(values={'tf': tfm}, stage_dv=Y, state_dv=y)['c']
    J, cols = blocks.dense(i)
    dcdtf = J[:, cols['h'][0]:cols['h'][0] + cols['h'][1]].sum(axis=1) / (NT - 1)
    fd = np.concatenate([np_((cp['stage_c'] - cm['stage_c'])[..., i]).reshape(-1),
                         np_((cp['state_c'] - cm['state_c'])[..., i]).reshape(-1)]) / (2 * e)
    assert np.allclose(dcdtf, fd, rtol=1e-6, atol=1e-8)


def test_compact_node_jacobians_equal_the_dense_ones_and_follow_the_declared_patterns():
    """oz_rhs_jac_eval with compact = 1 writes only the entries oz_rhs_jac_pattern marks; the dense output must be zero
    everywhere else (i.e. the functor's JP / JPP declarations are truthful) and equal on the marked entries."""
    from ozone_b200.glms import get_integration_method
    rng = np.random.default_rng(2)
    for rhs_name, d, P in (('lunar_ascent', 7, 3), ('van_der_pol', 3, 1), ('lotka_volterra', 2, 1)):
        spec = engine.SolveSpec(rhs_name=rhs_name, method=get_integration_method('RK4'), N=50, n=3)
        S = spec.info['S']
        y = torch.tensor(rng.uniform(0.5, 2.0, (6, d, 50)), dtype=torch.float64, device='cuda')
        p = torch.tensor(rng.uniform(-1.0, 1.0, (6, P, 50)), dtype=torch.float64, device='cuda')
        th = torch.tensor(rng.uniform(0.4, 0.6, (S, 50)), dtype=torch.float64, device='cuda') if S else None
        jp, jpp = engine.rhs_jac_pattern(spec)
        f0, Jy, Jp = engine.rhs_jac_eval(spec, y, p, th)
        f1, Jyc, Jpc = engine.rhs_jac_eval(spec, y, p, th, compact=True)
        assert torch.equal(f0, f1)
        assert tuple(Jyc.shape) == (6, int(jp.sum()), 50) and tuple(Jpc.shape) == (6, int(jpp.sum()), 50)
        iy, ip = np.argwhere(jp), np.argwhere(jpp)
        assert torch.equal(Jy[:, iy[:, 0], iy[:, 1], :], Jyc) and torch.equal(Jp[:, ip[:, 0], ip[:, 1], :], Jpc)
        mask_y = torch.tensor(~jp, device='cuda')
        mask_p = torch.tensor(~jpp, device='cuda')
        assert float(Jy[:, mask_y, :].abs().max()) == 0.0 if mask_y.any() else True
        assert float(Jp[:, mask_p, :].abs().max()) == 0.0 if mask_p.any() else True
    jp, jpp = engine.rhs_jac_pattern(engine.SolveSpec(rhs_name='lunar_ascent', method=get_integration_method('RK4'), N=1, n=1))
    assert int(jp.sum()) == 9 and int(jpp.sum()) == 3
