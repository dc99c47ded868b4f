"""oz_tm_adjoint_static: the reverse sweep for callers that cannot allocate device memory.

Some users of the library cannot give it a workspace: code running as a plug-in inside an
application that owns the GPU and forbids allocation in callbacks, control loops that cannot
afford cudaMalloc's synchronisation, or minimal FFI stubs that only forward the pointers they
were handed.  For them the library carries its own fixed 4 MiB scratch.  This test drives that
entry point through the C ABI exactly as such a caller would and checks that it returns the same
gradients as the regular, workspace-taking oz_tm_adjoint -- twice in a row, because the built-in
scratch is shared and keeps the previous call's contents.

The file name sorts last on purpose: a run that stops at the first failure has then already run
every other test.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from tests import cases
from tests.gpu_util import make_spec, dev_inputs, np_, relerr

pytestmark = pytest.mark.gpu


def _problem(case, method, grads):
    """Forward solve through the ABI, then a descriptor ready for either adjoint entry point."""
    from ozone_b200 import engine, _native as nat
    spec = make_spec(case, method)
    y0, h, p, th = dev_inputs(case, spec)
    fwd = engine.tm_forward(spec, y0, h, p, th)
    rng = np.random.default_rng(11)
    t = lambda a: torch.tensor(a, dtype=spec.dtype, device='cuda')
    cot = dict(hist_bar=t(rng.standard_normal(tuple(fwd['hist'].shape))),
               final_bar=t(rng.standard_normal(tuple(fwd['final'].shape))),
               prof_bar=t(rng.standard_normal(tuple(fwd['prof'].shape))) if fwd['prof'] is not None else None,
               field_bar=t(rng.standard_normal(tuple(fwd['field'].shape))) if fwd['field'] is not None else None)
    pb = engine.make_problem(spec)
    keep = engine._set_inputs(pb, spec, y0, h, p, th)
    pb.hist = fwd['hist'].data_ptr()
    for k, v in cot.items():
        setattr(pb, k, None if v is None else v.data_ptr())
    shp = engine.input_shapes(spec)
    out = {k: torch.full(shp[k], float('nan'), dtype=spec.dtype, device='cuda') for k in grads}
    for k in grads:
        setattr(pb, k + '_bar', out[k].data_ptr())
    return nat, pb, out, (keep, fwd, cot)


@pytest.mark.parametrize('name,method', [('lotka_volterra', 'RK4'), ('van_der_pol', 'GaussLegendre4')])
def test_static_scratch_adjoint_matches_workspace_adjoint_twice(name, method):
    if name == 'lotka_volterra':      # shared a(t) and shared h: both gradients go through the scratch
        case, grads = cases.lotka_volterra_inputs(N=3000, nt=41), ('y0', 'h', 'p', 'th')
    else:
        case, grads = cases.van_der_pol_inputs(N=1500, nt=21), ('y0', 'h', 'p')
    nat, pb, out, keep = _problem(case, method, grads)

    # reference result: the existing entry point with a caller-allocated workspace
    need = nat.lib.oz_tm_adjoint_workspace_bytes(C.byref(pb))
    assert 0 < need <= nat.lib.oz_static_scratch_bytes()
    ws = torch.empty(need, dtype=torch.uint8, device='cuda')
    pb.ws, pb.ws_bytes = ws.data_ptr(), need
    nat.check(nat.lib.oz_tm_adjoint(C.byref(pb)), 'oz_tm_adjoint')
    torch.cuda.synchronize()
    expect = {k: np_(v) for k, v in out.items()}
    assert all(np.isfinite(v).all() for v in expect.values())

    # the caller that cannot allocate passes no workspace at all
    pb.ws, pb.ws_bytes = None, 0
    for call in range(2):
        for v in out.values():
            v.fill_(float('nan'))
        nat.check(nat.lib.oz_tm_adjoint_static(C.byref(pb)), 'oz_tm_adjoint_static')
        torch.cuda.synchronize()
        for k, v in out.items():
            assert relerr(np_(v), expect[k]) < 1e-10, (call, k)


@pytest.mark.parametrize('name,method', [('lotka_volterra', 'RK4'), ('van_der_pol', 'GaussLegendre4')])
def test_static_scratch_adjoint_with_checkpoints_replays_segments_in_the_builtin_scratch(name, method):
    """TimeMarchingCheckpoints without a stored history: besides the partial sums, the built-in scratch then also
    holds the states of the segment being replayed (num_inner x d x N values)."""
    from ozone_b200 import engine, _native as nat
    if name == 'lotka_volterra':
        case, grads = cases.lotka_volterra_inputs(N=2000, nt=41), ('y0', 'h', 'p', 'th')
    else:
        case, grads = cases.van_der_pol_inputs(N=1100, nt=21), ('y0', 'h', 'p')
    ni = 5
    spec = make_spec(case, method, num_inner=ni, store_history=False)
    y0, h, p, th = dev_inputs(case, spec)
    fwd = engine.tm_forward(spec, y0, h, p, th, keep_history=False)
    assert fwd['hist'] is None and fwd['ckpt'] is not None
    rng = np.random.default_rng(12)
    t = lambda a: torch.tensor(a, dtype=spec.dtype, device='cuda')
    final_bar = t(rng.standard_normal(tuple(fwd['final'].shape)))
    prof_bar = t(rng.standard_normal(tuple(fwd['prof'].shape))) if fwd['prof'] is not None else None
    field_bar = t(rng.standard_normal(tuple(fwd['field'].shape))) if fwd['field'] is not None else None
    pb = engine.make_problem(spec)
    keep = engine._set_inputs(pb, spec, y0, h, p, th)
    pb.ckpt, pb.final_bar = fwd['ckpt'].data_ptr(), final_bar.data_ptr()
    pb.prof_bar = None if prof_bar is None else prof_bar.data_ptr()
    pb.field_bar = None if field_bar is None else field_bar.data_ptr()
    shp = engine.input_shapes(spec)
    out = {k: torch.full(shp[k], float('nan'), dtype=spec.dtype, device='cuda') for k in grads}
    for k in grads:
        setattr(pb, k + '_bar', out[k].data_ptr())
    need = nat.lib.oz_tm_adjoint_workspace_bytes(C.byref(pb))
    assert ni * spec.d * spec.N * 8 <= need <= nat.lib.oz_static_scratch_bytes()       # the segment states are in there
    ws = torch.empty(need, dtype=torch.uint8, device='cuda')
    pb.ws, pb.ws_bytes = ws.data_ptr(), need
    nat.check(nat.lib.oz_tm_adjoint(C.byref(pb)), 'oz_tm_adjoint')
    torch.cuda.synchronize()
    expect = {k: np_(v) for k, v in out.items()}
    assert all(np.isfinite(v).all() for v in expect.values())
    pb.ws, pb.ws_bytes = None, 0
    for call in range(2):
        for v in out.values():
            v.fill_(float('nan'))
        nat.check(nat.lib.oz_tm_adjoint_static(C.byref(pb)), 'oz_tm_adjoint_static')
        torch.cuda.synchronize()
        for k, v in out.items():
            assert relerr(np_(v), expect[k]) < 1e-10, (call, k)
    del keep


def test_static_scratch_refuses_problems_that_do_not_fit():
    case = cases.lotka_volterra_inputs(N=64, nt=41)
    nat, pb, out, keep = _problem(case, 'RK4', ('h', 'p'))
    pb.N = 1 << 24            # 524,288 warps x 82 shared rows x 8 B is far beyond 4 MiB; refused before any launch
    rc = nat.lib.oz_tm_adjoint_static(C.byref(pb))
    assert rc < 0
    msg = nat.lib.oz_last_error().decode()
    assert 'scratch' in msg and str(nat.lib.oz_static_scratch_bytes()) in msg
    torch.cuda.synchronize()
