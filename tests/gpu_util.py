"""Helpers for the GPU parity tests: run one case through the C ABI (via ozone_b200.engine) and
through the oracle, return numpy arrays in the oracle's conventions."""
import numpy as np
import torch

from oracle import rhs as R, integrators as I, adjoint as Adj
from ozone_b200 import engine
from ozone_b200.glms import get_integration_method


def oracle_rhs(case):
    if case['rhs'] == 'diffusion_reaction':
        return R.DiffusionReaction(case['nx'])
    return R.REGISTRY[case['rhs']]()


def relerr(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    den = np.abs(b).max()
    return np.abs(a - b).max() / (den if den > 0 else 1.0)


def make_spec(case, method, dtype=torch.float64, num_inner=0, store_history=True, tol=1e-7, **kw):
    y0, h, p, th = case['y0'], case['h'], case['p'], case['th']
    N, n = y0.shape[1], h.shape[0]
    extra = {}
    if case['rhs'] == 'diffusion_reaction':
        nx = case['nx']
        extra = dict(d_override=nx, rhs_cfg=(np.pi / nx, 16.0, 2.0, 2.0, float(round(0.3 * nx)), float(round(0.7 * nx))))
    return engine.SolveSpec(
        rhs_name=case['rhs'], method=get_integration_method(method), N=N, n=n, dtype=dtype,
        h_shared=(h.ndim == 1), p_shared=(p is None or p.shape[-1] == 1),
        th_shared=(th is None or th.shape[-1] == 1), num_inner=num_inner, store_history=store_history,
        newton_tol=tol, **extra, **kw)


def dev_inputs(case, spec, requires_grad=False):
    dt = spec.dtype
    t = lambda a: torch.tensor(np.ascontiguousarray(a), dtype=dt, device='cuda', requires_grad=requires_grad)
    y0 = t(case['y0'])
    h = t(case['h'])
    p = th = None
    if spec.info['P']:
        p = t(case['p'][:, :, 0] if spec.p_shared else case['p'])
    if spec.info['S']:
        th = t(case['th'][:, 0] if spec.th_shared else case['th'])
    return y0, h, p, th


def np_(t):
    return None if t is None else t.detach().cpu().numpy().astype(np.float64)


def unshare(g, ref):
    """bring a device gradient of a shared input back to the oracle's broadcastable shape"""
    if g is None:
        return None
    g = np_(g)
    return g.reshape(ref.shape) if g.size == ref.size else g


def run_pair(case, method, dtype=torch.float64, cot_seed=0, tol=1e-7, num_inner=0, store_history=True, cache_stages=False):
    """Forward + adjoint through the engine and through the oracle with the same random cotangents."""
    spec = make_spec(case, method, dtype, num_inner=num_inner, store_history=store_history, tol=tol)
    y0, h, p, th = dev_inputs(case, spec)
    fwd = engine.tm_forward(spec, y0, h, p, th, keep_history=store_history, cache_stages=cache_stages)
    rhs = oracle_rhs(case)
    if num_inner:
        ref = I.time_marching_checkpoints(rhs, method, case['y0'], case['h'], case['p'], case['th'],
                                          num_checkpoints=len(case['h']) // num_inner, tol=tol)
    else:
        ref = I.time_marching(rhs, method, case['y0'], case['h'], case['p'], case['th'], tol=tol)
    rng = np.random.default_rng(cot_seed)
    nt, d, N = ref['history'].shape
    hb = rng.standard_normal((nt, d, N)) if store_history else None
    fb = rng.standard_normal((d, N))
    pb = rng.standard_normal(ref['profile'].shape) if rhs.NPROF else None
    qb = rng.standard_normal(ref['field'].shape) if rhs.NFIELD else None
    dt = spec.dtype
    t = lambda a: None if a is None else torch.tensor(a, dtype=dt, device='cuda')
    g = engine.tm_adjoint(spec, y0, h, p, th, hist=fwd['hist'], ckpt=fwd['ckpt'], hist_bar=t(hb), final_bar=t(fb),
                          prof_bar=t(pb), field_bar=t(qb), status=fwd['status'], stages=fwd.get('stages'))
    torch.cuda.synchronize()
    gref = Adj.tm_adjoint(rhs, method, case['y0'], case['h'], case['p'], case['th'], history=ref['history'],
                          hist_bar=hb, final_bar=fb, prof_bar=pb, field_bar=qb, tol=tol)
    return spec, fwd, ref, g, gref
