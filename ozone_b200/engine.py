"""Host-side plumbing between packed torch tensors and the C ABI.

Everything numeric happens in ``liboz_b200.so``; torch provides device memory, the current
stream and the autograd tape.  Packed layouts are the ones documented in
``include/ozone_b200.h`` (structure of arrays, trajectory index innermost).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field, replace
from typing import Optional

import torch

from . import _native as nat
from .glms import GLMMethod, MAX_STAGES


@dataclass
class SolveSpec:
    """Static description of one solve (everything that is not a differentiable tensor)."""
    rhs_name: str
    method: GLMMethod
    N: int
    n: int
    dtype: torch.dtype = torch.float64
    h_shared: bool = True
    p_shared: bool = False
    th_shared: bool = False
    num_inner: int = 0                 # > 0: TimeMarchingCheckpoints segment length
    store_history: bool = True
    newton_tol: float = 1e-7           # ozone/integrators/time_marching.py:220
    newton_max_iter: int = 100
    d_override: int = 0                # grid RHS: number of cells
    rhs_cfg: tuple = ()
    want_status: bool = True
    state_blocks: tuple = ()           # Picard: packed-component offsets of the states (block Gauss-Seidel order)
    grid_slab: tuple = ()              # grid RHS slab mode: (x_off, nx_global, own_lo, own_hi, flags)
    traced: object = None              # trace.TracedRHS: a Python callback compiled to a plugin functor
    info: dict = field(default_factory=dict)

    def __post_init__(self):
        if self.traced is not None:
            t = self.traced
            self.rhs_id = None          # resolved per (stages, implicit, dtype, approach) in make_problem
            if t.kind == 'grid':           # trace_grid.TracedGridRHS: the cells come with the problem, coefficient fields as th
                self.info = t.spec_info(self.d_override)
            else:
                self.info = dict(d=t.D, P=t.P, S=t.S, nprof=len(t.prof_exprs), nfield=len(t.field_exprs), kind=0)
        else:
            self.rhs_id = nat.RHS_IDS[self.rhs_name]
        if not self.info:
            self.info = nat.rhs_info(self.rhs_id)
        if self.info['kind'] == 1:
            self.info = dict(self.info, d=int(self.d_override))
        if self.dtype not in (torch.float64, torch.float32):
            raise TypeError(f'unsupported dtype {self.dtype}; the kernels compute in float64 or float32')

    @property
    def d(self):
        return self.info['d']

    @property
    def nt(self):
        return self.n + 1


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None or t.numel() == 0 else t.data_ptr()


def _chk(t: Optional[torch.Tensor], spec: SolveSpec, shape, name):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f'{name} must live on a CUDA device (ozone_b200 has no CPU path)')
    if t.dtype != spec.dtype:
        raise TypeError(f'{name} has dtype {t.dtype}, expected {spec.dtype}')
    if tuple(t.shape) != tuple(shape):
        raise ValueError(f'{name} has shape {tuple(t.shape)}, expected {tuple(shape)}')
    return t.contiguous()


KIND_TM, KIND_PICARD, KIND_COLLOC = 0, 1, 2


def make_problem(spec: SolveSpec, kind: int = KIND_TM) -> nat.OzProblem:
    pb = nat.OzProblem()
    m = spec.method
    s = m.num_stages
    if s > MAX_STAGES:
        raise ValueError(f'{s} stages exceed the device tableau storage ({MAX_STAGES})')
    rhs_id = spec.rhs_id
    if spec.traced is not None:         # compiled on first use for this (stages, implicit, dtype, approach)
        rhs_id = spec.traced.plugin_id(s, not m.explicit if kind == KIND_TM else False, spec.dtype == torch.float64, kind)
    pb.rhs_id, pb.dtype = rhs_id, (nat.OZ_F64 if spec.dtype == torch.float64 else nat.OZ_F32)
    pb.num_stages, pb.is_explicit = s, int(m.explicit)
    for i in range(s):
        for j in range(s):
            pb.A[i * MAX_STAGES + j] = float(m.A[i, j])
        pb.B[i] = float(m.B[0, i])
        pb.C[i] = float(m.C[i, 0])
    pb.N, pb.n, pb.num_inner = spec.N, spec.n, spec.num_inner
    pb.h_shared, pb.p_shared, pb.th_shared = int(spec.h_shared), int(spec.p_shared), int(spec.th_shared)
    pb.d_override = int(spec.d_override)
    pb.newton_tol, pb.newton_max_iter = float(spec.newton_tol), int(spec.newton_max_iter)
    for i, v in enumerate(spec.rhs_cfg):
        pb.rhs_cfg[i] = float(v)
    if spec.grid_slab:
        pb.grid_x_off, pb.grid_nx_global, pb.grid_own_lo, pb.grid_own_hi, pb.grid_flags = (int(v) for v in spec.grid_slab)
    if spec.state_blocks:
        pb.num_state_blocks = len(spec.state_blocks) - 1
        for i, v in enumerate(spec.state_blocks):
            pb.state_block_lo[i] = int(v)
    pb.stream = torch.cuda.current_stream().cuda_stream
    return pb


def input_shapes(spec: SolveSpec):
    d, P, S, N, n = spec.d, spec.info['P'], spec.info['S'], spec.N, spec.n
    return dict(
        y0=(d, N),
        h=(n,) if spec.h_shared else (n, N),
        p=(n + 1, P) if spec.p_shared else (n + 1, P, N),
        th=(S,) if spec.th_shared else (S, N),
    )


def _set_inputs(pb, spec, y0, h, p, th):
    shp = input_shapes(spec)
    y0 = _chk(y0, spec, shp['y0'], 'y0')
    h = _chk(h, spec, shp['h'], 'h')
    p = _chk(p, spec, shp['p'], 'p') if spec.info['P'] else None
    th = _chk(th, spec, shp['th'], 'th') if spec.info['S'] else None
    pb.y0, pb.h, pb.p, pb.th = _ptr(y0), _ptr(h), _ptr(p), _ptr(th)
    return y0, h, p, th


def tm_forward(spec: SolveSpec, y0, h, p=None, th=None, keep_history=None, cache_stages=False):
    """oz_tm_forward.  Returns dict(hist, ckpt, final, prof, field, status, iters); absent outputs are None."""
    dev = y0.device
    pb = make_problem(spec)
    keep = _set_inputs(pb, spec, y0, h, p, th)
    d, N, nt = spec.d, spec.N, spec.nt
    keep_history = spec.store_history if keep_history is None else keep_history
    new = lambda *shape: torch.empty(shape, dtype=spec.dtype, device=dev)
    out = dict(hist=None, ckpt=None, final=new(d, N), prof=None, field=None, status=None, iters=None)
    if keep_history:
        out['hist'] = new(nt, d, N)
    if spec.num_inner > 0:
        out['ckpt'] = new(spec.n // spec.num_inner + 1, d, N)
    if spec.info['nprof']:
        out['prof'] = new(nt, spec.info['nprof'], N)
    if spec.info['nfield']:
        out['field'] = new(spec.info['nfield'], N)
    if spec.want_status:
        out['status'] = torch.zeros(N, dtype=torch.int32, device=dev)
        out['iters'] = torch.zeros(N, dtype=torch.int32, device=dev)
    pb.hist, pb.ckpt, pb.yfinal = _ptr(out['hist']), _ptr(out['ckpt']), _ptr(out['final'])
    pb.prof, pb.field = _ptr(out['prof']), _ptr(out['field'])
    pb.status, pb.iters = _ptr(out['status']), _ptr(out['iters'])
    out['stages'] = None
    if cache_stages and not spec.method.explicit and spec.info['kind'] == 0:
        out['stages'] = new(spec.n, spec.method.num_stages, d, N)
        pb.stages = _ptr(out['stages'])
    ws = _forward_workspace(pb, dev)
    nat.check(nat.lib.oz_tm_forward(C.byref(pb)), 'oz_tm_forward')
    del keep, ws
    return out


def _forward_workspace(pb, dev):
    """Grid right-hand sides need scratch for the forward sweep too (ping-pong state + partial sums)."""
    ws_bytes = nat.lib.oz_tm_forward_workspace_bytes(C.byref(pb))
    if ws_bytes == C.c_size_t(-1).value:
        raise nat.OzError(f'oz_tm_forward_workspace_bytes: {nat.lib.oz_last_error().decode()}')
    if ws_bytes == 0:
        return None
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    pb.ws, pb.ws_bytes = ws.data_ptr(), ws_bytes
    return ws


def tm_adjoint(spec: SolveSpec, y0, h, p, th, hist=None, ckpt=None, hist_bar=None, final_bar=None,
               prof_bar=None, field_bar=None, need=('y0', 'h', 'p', 'th'), status=None, stages=None):
    """oz_tm_adjoint.  Returns dict(y0, h, p, th) of gradients shaped like the inputs."""
    dev = y0.device
    pb = make_problem(spec)
    keep = _set_inputs(pb, spec, y0, h, p, th)
    d, N, nt = spec.d, spec.N, spec.nt
    shp = input_shapes(spec)
    hist = _chk(hist, spec, (nt, d, N), 'hist')
    ckpt = _chk(ckpt, spec, (spec.n // spec.num_inner + 1, d, N), 'ckpt') if spec.num_inner > 0 else None
    hist_bar = _chk(hist_bar, spec, (nt, d, N), 'hist_bar')
    final_bar = _chk(final_bar, spec, (d, N), 'final_bar')
    prof_bar = _chk(prof_bar, spec, (nt, spec.info['nprof'], N), 'prof_bar') if spec.info['nprof'] else None
    field_bar = _chk(field_bar, spec, (spec.info['nfield'], N), 'field_bar') if spec.info['nfield'] else None
    pb.hist, pb.ckpt = _ptr(hist), _ptr(ckpt)
    if stages is not None:
        pb.stages = _ptr(_chk(stages, spec, (spec.n, spec.method.num_stages, d, N), 'stages'))
    pb.hist_bar, pb.final_bar, pb.prof_bar, pb.field_bar = map(_ptr, (hist_bar, final_bar, prof_bar, field_bar))
    grads = {}
    for k in ('y0', 'h', 'p', 'th'):
        size_ok = all(s > 0 for s in shp[k])
        grads[k] = torch.zeros(shp[k], dtype=spec.dtype, device=dev) if (k in need and size_ok) else None
    pb.y0_bar, pb.h_bar, pb.p_bar, pb.th_bar = (_ptr(grads[k]) for k in ('y0', 'h', 'p', 'th'))
    pb.status = _ptr(status)
    ws_bytes = nat.lib.oz_tm_adjoint_workspace_bytes(C.byref(pb))
    if ws_bytes == C.c_size_t(-1).value:
        raise nat.OzError(f'oz_tm_adjoint_workspace_bytes: {nat.lib.oz_last_error().decode()}')
    ws = torch.empty(max(ws_bytes, 8), dtype=torch.uint8, device=dev)
    pb.ws, pb.ws_bytes = ws.data_ptr(), ws_bytes
    nat.check(nat.lib.oz_tm_adjoint(C.byref(pb)), 'oz_tm_adjoint')
    del keep
    return grads


STAGE_CACHE_LIMIT = 8 << 30      # bytes; beyond this the reverse sweep re-solves the stages instead


class _TimeMarch(torch.autograd.Function):
    """Solve + reverse sweep on the autograd tape.  Inputs are packed; outputs are
    (hist | empty, final, prof | empty, field | empty, status, iters)."""

    @staticmethod
    def forward(ctx, spec: SolveSpec, y0, h, p, th):
        needs = [t is not None and t.requires_grad for t in (y0, h, p, th)]
        any_grad = any(needs)
        keep_hist = spec.store_history or (any_grad and spec.num_inner == 0)
        # implicit methods: keep the converged stages so the reverse sweep does not repeat the Newton solves
        stage_bytes = spec.n * spec.method.num_stages * spec.d * spec.N * (8 if spec.dtype == torch.float64 else 4)
        cache = any_grad and keep_hist and spec.num_inner == 0 and not spec.method.explicit and stage_bytes <= STAGE_CACHE_LIMIT
        out = tm_forward(spec, y0, h, p, th, keep_history=keep_hist, cache_stages=cache)
        ctx.spec, ctx.needs = spec, needs
        ctx.has = dict(hist=out['hist'] is not None, prof=out['prof'] is not None, field=out['field'] is not None)
        ctx.status = out['status']
        saved = [y0, h, p, th, out['hist'], out['ckpt'], out['stages']]
        ctx.present = [t is not None for t in saved]
        ctx.save_for_backward(*[t for t in saved if t is not None])
        e = lambda: torch.empty(0, dtype=spec.dtype, device=y0.device)
        hist = out['hist'] if (spec.store_history and out['hist'] is not None) else e()
        ret = (hist, out['final'], out['prof'] if out['prof'] is not None else e(),
               out['field'] if out['field'] is not None else e(),
               out['status'] if out['status'] is not None else torch.empty(0, dtype=torch.int32, device=y0.device),
               out['iters'] if out['iters'] is not None else torch.empty(0, dtype=torch.int32, device=y0.device))
        ctx.mark_non_differentiable(ret[4], ret[5])
        ctx.hist_returned = hist.numel() > 0
        return ret

    @staticmethod
    def backward(ctx, g_hist, g_final, g_prof, g_field, _gs, _gi):
        spec = ctx.spec
        it = iter(ctx.saved_tensors)
        y0, h, p, th, hist, ckpt, stages = [next(it) if present else None for present in ctx.present]
        need = tuple(k for k, flag in zip(('y0', 'h', 'p', 'th'), ctx.needs) if flag)
        fix = lambda g, ok: g.contiguous() if (ok and g is not None and g.numel() > 0) else None
        grads = tm_adjoint(spec, y0, h, p, th, hist=hist, ckpt=ckpt, stages=stages,
                           hist_bar=fix(g_hist, ctx.hist_returned), final_bar=fix(g_final, True),
                           prof_bar=fix(g_prof, ctx.has['prof']), field_bar=fix(g_field, ctx.has['field']),
                           need=need, status=ctx.status)
        return (None,) + tuple(grads[k] if flag else None for k, flag in zip(('y0', 'h', 'p', 'th'), ctx.needs))


def time_march(spec: SolveSpec, y0, h, p=None, th=None):
    """Differentiable packed solve: returns dict(hist, final, prof, field, status, iters)."""
    hist, final, prof, fld, status, iters = _TimeMarch.apply(spec, y0, h, p, th)
    none_if_empty = lambda t: t if t.numel() > 0 else None
    return dict(hist=none_if_empty(hist), final=final, prof=none_if_empty(prof), field=none_if_empty(fld),
                status=none_if_empty(status), iters=none_if_empty(iters))


def rhs_eval(spec: SolveSpec, y, p=None, th=None):
    """oz_rhs_eval: f and the outputs at `nodes` points; y (d, nodes), p (P, nodes)."""
    pb = make_problem(spec)
    nodes = y.shape[1]
    info = spec.info
    dev = y.device
    y = y.contiguous()
    p = p.contiguous() if (p is not None and info['P']) else None
    th = th.contiguous() if (th is not None and info['S']) else None
    pb.th = _ptr(th)
    pb.th_shared = int(th is None or th.dim() == 1)
    dy = torch.empty_like(y)
    prof = torch.empty((info['nprof'], nodes), dtype=spec.dtype, device=dev) if info['nprof'] else None
    fld = torch.empty((info['nfield'], nodes), dtype=spec.dtype, device=dev) if info['nfield'] else None
    nat.check(nat.lib.oz_rhs_eval(C.byref(pb), nodes, _ptr(y), _ptr(p), _ptr(dy), _ptr(prof), _ptr(fld)), 'oz_rhs_eval')
    return dy, prof, fld


GRID_STEPS_PER_LAUNCH = 4      # OZ_PDE_FUSE in csrc/oz_pde.cu: forward grid steps fused into one launch
GRID_REVERSE_STEPS_PER_LAUNCH = 2   # OZ_PDE_FUSE_ADJ


def grid_fusion(nx):
    """(forward, reverse) steps fused per launch for a grid of nx cells: mirrors pde_layout() in csrc/oz_pde.cu
    (launch counting only)."""
    if nx <= 160000:
        return 8, 4
    return (16, 8) if nx <= 600000 else (GRID_STEPS_PER_LAUNCH, GRID_REVERSE_STEPS_PER_LAUNCH)


class TmPlan:
    """A solve with every buffer allocated once: repeated forward()/adjoint() calls only enqueue
    kernels (no allocation, no host sync).  This is what a caller inside an optimisation loop
    uses; ``bench.py`` times it."""

    def __init__(self, spec: SolveSpec, device='cuda', grads=('y0', 'h', 'p', 'th'), cotangents=('final',),
                 cache_stages=True):
        self.spec = spec
        dev = torch.device(device)
        d, N, n, nt = spec.d, spec.N, spec.n, spec.nt
        info = spec.info
        new = lambda *shape: torch.zeros(shape, dtype=spec.dtype, device=dev)
        shp = input_shapes(spec)
        self.y0, self.h = new(*shp['y0']), new(*shp['h'])
        self.p = new(*shp['p']) if info['P'] else None
        self.th = new(*shp['th']) if info['S'] else None
        self.hist = new(nt, d, N) if (spec.store_history or spec.num_inner == 0) else None
        self.ckpt = new(n // spec.num_inner + 1, d, N) if spec.num_inner > 0 else None
        self.final = new(d, N)
        self.prof = new(nt, info['nprof'], N) if info['nprof'] else None
        self.field = new(info['nfield'], N) if info['nfield'] else None
        self.status = torch.zeros(N, dtype=torch.int32, device=dev)
        self.iters = torch.zeros(N, dtype=torch.int32, device=dev)
        self.hist_bar = new(nt, d, N) if 'hist' in cotangents else None
        self.final_bar = new(d, N) if 'final' in cotangents else None
        self.prof_bar = new(nt, info['nprof'], N) if ('prof' in cotangents and info['nprof']) else None
        self.field_bar = new(info['nfield'], N) if ('field' in cotangents and info['nfield']) else None
        self.grad = {k: (new(*shp[k]) if (k in grads and all(s > 0 for s in shp[k])) else None)
                     for k in ('y0', 'h', 'p', 'th')}
        pb = make_problem(spec)
        pb.y0, pb.h, pb.p, pb.th = map(_ptr, (self.y0, self.h, self.p, self.th))
        pb.hist, pb.ckpt, pb.yfinal = _ptr(self.hist), _ptr(self.ckpt), _ptr(self.final)
        pb.prof, pb.field = _ptr(self.prof), _ptr(self.field)
        pb.status, pb.iters = _ptr(self.status), _ptr(self.iters)
        pb.hist_bar, pb.final_bar = _ptr(self.hist_bar), _ptr(self.final_bar)
        pb.prof_bar, pb.field_bar = _ptr(self.prof_bar), _ptr(self.field_bar)
        pb.y0_bar, pb.h_bar, pb.p_bar, pb.th_bar = (_ptr(self.grad[k]) for k in ('y0', 'h', 'p', 'th'))
        self.stages = None
        if cache_stages and not spec.method.explicit and info['kind'] == 0 and self.hist is not None:
            self.stages = new(n, spec.method.num_stages, d, N)
            pb.stages = _ptr(self.stages)
        ws_bytes = nat.lib.oz_tm_adjoint_workspace_bytes(C.byref(pb))
        if ws_bytes == C.c_size_t(-1).value:
            raise nat.OzError(f'oz_tm_adjoint_workspace_bytes: {nat.lib.oz_last_error().decode()}')
        fwd_bytes = nat.lib.oz_tm_forward_workspace_bytes(C.byref(pb))
        if fwd_bytes == C.c_size_t(-1).value:
            raise nat.OzError(f'oz_tm_forward_workspace_bytes: {nat.lib.oz_last_error().decode()}')
        ws_bytes = max(ws_bytes, fwd_bytes)           # forward and adjoint never run concurrently: one buffer
        self.ws = torch.empty(max(ws_bytes, 8), dtype=torch.uint8, device=dev)
        pb.ws, pb.ws_bytes = self.ws.data_ptr(), ws_bytes
        self.pb = pb
        # kernels of ours enqueued by one forward() / one adjoint() call
        shared_rows = sum(1 for k, flag in (('h', spec.h_shared), ('p', spec.p_shared), ('th', spec.th_shared))
                          if flag and self.grad[k] is not None)
        self.launches_forward, self.launches_adjoint = 1, 1 + shared_rows
        if spec.info['kind'] == 1:      # grid: fused multi-step launches (+ sum-of-squares, field / terminal, replay, gradient reducers)
            (K, KR), ni = grid_fusion(spec.d), spec.num_inner
            groups = lambda steps: -(-steps // K)          # launches of the fused forward kernel for a run of steps
            nseg = (n // ni) if (ni > 0 and self.hist is None) else 0
            self.launches_forward = ((n // ni) * groups(ni) if ni > 0 else groups(n)) + 2
            rgroups = lambda steps: -(-steps // KR)
            self.launches_adjoint = 1 + (nseg * (groups(ni - 1) + rgroups(ni)) if nseg else rgroups(n)) + 1

    def forward(self):
        self.pb.stream = torch.cuda.current_stream().cuda_stream
        nat.check(nat.lib.oz_tm_forward(C.byref(self.pb)), 'oz_tm_forward')

    def adjoint(self):
        self.pb.stream = torch.cuda.current_stream().cuda_stream
        nat.check(nat.lib.oz_tm_adjoint(C.byref(self.pb)), 'oz_tm_adjoint')


# ------------------------------------------------------------------------------------------------
def picard_forward(spec: SolveSpec, y0, h, p=None, th=None, guess=None, tol=1e-10, max_sweeps=2000):
    """oz_picard_forward: all (step, stage) nodes relaxed at once; returns the same dict as tm_forward."""
    dev = y0.device
    pb = make_problem(spec, KIND_PICARD)
    keep = _set_inputs(pb, spec, y0, h, p, th)
    d, N, nt = spec.d, spec.N, spec.nt
    guess = _chk(guess, spec, (spec.n, d, N), 'guess')
    new = lambda *shape: torch.empty(shape, dtype=spec.dtype, device=dev)
    out = dict(hist=new(nt, d, N), ckpt=None, final=new(d, N), prof=None, field=None,
               status=torch.zeros(N, dtype=torch.int32, device=dev), iters=torch.zeros(N, dtype=torch.int32, device=dev))
    if spec.info['nprof']:
        out['prof'] = new(nt, spec.info['nprof'], N)
    if spec.info['nfield']:
        out['field'] = new(spec.info['nfield'], N)
    pb.hist, pb.yfinal, pb.prof, pb.field = _ptr(out['hist']), _ptr(out['final']), _ptr(out['prof']), _ptr(out['field'])
    pb.status, pb.iters = _ptr(out['status']), _ptr(out['iters'])
    nat.check(nat.lib.oz_picard_forward(C.byref(pb), _ptr(guess), float(tol), int(max_sweeps)), 'oz_picard_forward')
    del keep
    return out


class _Picard(torch.autograd.Function):
    """Picard solve on the tape.  The converged iterate satisfies the same discrete equations as
    time marching, so its reverse sweep is the time-marching adjoint evaluated on the solved history
    (implicit-function theorem: no need to differentiate through the sweeps)."""

    @staticmethod
    def forward(ctx, spec, guess, tol, max_sweeps, y0, h, p, th):
        out = picard_forward(spec, y0, h, p, th, guess=guess, tol=tol, max_sweeps=max_sweeps)
        ctx.spec = replace(spec, num_inner=0)
        ctx.needs = [t is not None and t.requires_grad for t in (y0, h, p, th)]
        saved = [y0, h, p, th, out['hist']]
        ctx.present = [t is not None for t in saved]
        ctx.save_for_backward(*[t for t in saved if t is not None])
        ctx.has = dict(prof=out['prof'] is not None, field=out['field'] is not None)
        e = lambda: torch.empty(0, dtype=spec.dtype, device=y0.device)
        ret = (out['hist'], out['final'], out['prof'] if out['prof'] is not None else e(),
               out['field'] if out['field'] is not None else e(), out['status'], out['iters'])
        ctx.mark_non_differentiable(ret[4], ret[5])
        return ret

    @staticmethod
    def backward(ctx, g_hist, g_final, g_prof, g_field, _gs, _gi):
        it = iter(ctx.saved_tensors)
        y0, h, p, th, hist = [next(it) if present else None for present in ctx.present]
        need = tuple(k for k, flag in zip(('y0', 'h', 'p', 'th'), ctx.needs) if flag)
        fix = lambda g, ok: g.contiguous() if (ok and g is not None and g.numel() > 0) else None
        grads = tm_adjoint(ctx.spec, y0, h, p, th, hist=hist, hist_bar=fix(g_hist, True), final_bar=fix(g_final, True),
                           prof_bar=fix(g_prof, ctx.has['prof']), field_bar=fix(g_field, ctx.has['field']), need=need)
        return (None, None, None, None) + tuple(grads[k] if flag else None
                                                for k, flag in zip(('y0', 'h', 'p', 'th'), ctx.needs))


def picard(spec: SolveSpec, y0, h, p=None, th=None, guess=None, tol=1e-10, max_sweeps=2000):
    hist, final, prof, fld, status, iters = _Picard.apply(spec, guess, tol, max_sweeps, y0, h, p, th)
    none_if_empty = lambda t: t if t.numel() > 0 else None
    return dict(hist=hist if spec.store_history else None, final=final, prof=none_if_empty(prof),
                field=none_if_empty(fld), status=status, iters=iters)


# ------------------------------------------------------------------------------------------------
def colloc_residual(spec: SolveSpec, y0, h, p, th, Y, y):
    dev = y0.device
    pb = make_problem(spec, KIND_COLLOC)
    keep = _set_inputs(pb, spec, y0, h, p, th)
    n, s, d, N = spec.n, spec.method.num_stages, spec.d, spec.N
    Y = _chk(Y, spec, (n, s, d, N), 'stage design variables')
    y = _chk(y, spec, (n, d, N), 'state design variables')
    c1 = torch.empty((n, s, d, N), dtype=spec.dtype, device=dev)
    c2 = torch.empty((n, d, N), dtype=spec.dtype, device=dev)
    nat.check(nat.lib.oz_colloc_residual(C.byref(pb), _ptr(Y), _ptr(y), _ptr(c1), _ptr(c2)), 'oz_colloc_residual')
    del keep
    return c1, c2


def colloc_vjp(spec: SolveSpec, y0, h, p, th, Y, y, mu1, mu2, need=('y0', 'h', 'p', 'th')):
    """oz_colloc_vjp: gradient of <mu1, c1> + <mu2, c2> w.r.t. (Y, y, y0, h, p, th)."""
    dev = y0.device
    pb = make_problem(spec, KIND_COLLOC)
    keep = _set_inputs(pb, spec, y0, h, p, th)
    n, s, d, N = spec.n, spec.method.num_stages, spec.d, spec.N
    Y = _chk(Y, spec, (n, s, d, N), 'stage design variables')
    y = _chk(y, spec, (n, d, N), 'state design variables')
    mu1 = _chk(mu1, spec, (n, s, d, N), 'mu1')
    mu2 = _chk(mu2, spec, (n, d, N), 'mu2')
    shp = input_shapes(spec)
    grads = {k: (torch.zeros(shp[k], dtype=spec.dtype, device=dev) if (k in need and all(v > 0 for v in shp[k])) else None)
             for k in ('y0', 'h', 'p', 'th')}
    pb.y0_bar, pb.h_bar, pb.p_bar, pb.th_bar = (_ptr(grads[k]) for k in ('y0', 'h', 'p', 'th'))
    Yb, yb = torch.empty_like(Y), torch.empty_like(y)
    ws_bytes = nat.lib.oz_colloc_vjp_workspace_bytes(C.byref(pb))
    if ws_bytes == C.c_size_t(-1).value:
        raise nat.OzError(f'oz_colloc_vjp_workspace_bytes: {nat.lib.oz_last_error().decode()}')
    ws = torch.empty(max(ws_bytes, 8), dtype=torch.uint8, device=dev)
    pb.ws, pb.ws_bytes = ws.data_ptr(), ws_bytes
    nat.check(nat.lib.oz_colloc_vjp(C.byref(pb), _ptr(Y), _ptr(y), _ptr(mu1), _ptr(mu2), _ptr(Yb), _ptr(yb)),
              'oz_colloc_vjp')
    del keep
    return dict(grads, Y=Yb, y=yb)


class _Collocation(torch.autograd.Function):
    @staticmethod
    def forward(ctx, spec, y0, h, p, th, Y, y):
        c1, c2 = colloc_residual(spec, y0, h, p, th, Y, y)
        ctx.spec = spec
        ctx.needs = [t is not None and t.requires_grad for t in (y0, h, p, th, Y, y)]
        saved = [y0, h, p, th, Y, y]
        ctx.present = [t is not None for t in saved]
        ctx.save_for_backward(*[t for t in saved if t is not None])
        return c1, c2

    @staticmethod
    def backward(ctx, mu1, mu2):
        it = iter(ctx.saved_tensors)
        y0, h, p, th, Y, y = [next(it) if present else None for present in ctx.present]
        need = tuple(k for k, flag in zip(('y0', 'h', 'p', 'th'), ctx.needs[:4]) if flag)
        g = colloc_vjp(ctx.spec, y0, h, p, th, Y, y, mu1.contiguous(), mu2.contiguous(), need=need)
        order = ('y0', 'h', 'p', 'th', 'Y', 'y')
        return (None,) + tuple(g[k] if flag else None for k, flag in zip(order, ctx.needs))


def collocation(spec: SolveSpec, y0, h, p, th, Y, y):
    """Differentiable constraints (c1, c2) of ozone/integrators/collocation.py:48-50."""
    return _Collocation.apply(spec, y0, h, p, th, Y, y)


def _node_layout(spec: SolveSpec, hist, p, th):
    """(y, p, th) of a history as oz_rhs_eval wants them: every (time point, trajectory) pair is one node."""
    nt, d, N = hist.shape
    info = spec.info
    y = hist.permute(1, 0, 2).reshape(d, nt * N).contiguous()
    pp = None
    if info['P']:
        pp = (p.unsqueeze(-1).expand(nt, info['P'], N) if spec.p_shared else p).permute(1, 0, 2).reshape(info['P'], nt * N).contiguous()
    tt = None
    if info['S']:
        tt = th.contiguous() if spec.th_shared else th.unsqueeze(1).expand(info['S'], nt, N).reshape(info['S'], nt * N).contiguous()
    return y, pp, tt


def rhs_out_vjp(spec: SolveSpec, y, p, th, vprof, vfield):
    """oz_rhs_out_vjp: per-node gradients (y_bar (d, nodes), p_bar (P, nodes) | None, th_bar (S, nodes) | None)."""
    pb = make_problem(spec)
    info, nodes, dev = spec.info, y.shape[1], y.device
    pb.th = _ptr(th)
    pb.th_shared = int(th is None or th.dim() == 1)
    new = lambda rows: torch.empty((rows, nodes), dtype=spec.dtype, device=dev)
    yb = new(info['d'])
    pbar = new(info['P']) if info['P'] else None
    tbar = new(info['S']) if info['S'] else None
    nat.check(nat.lib.oz_rhs_out_vjp(C.byref(pb), nodes, _ptr(y), _ptr(p), _ptr(vprof), _ptr(vfield), _ptr(yb), _ptr(pbar),
                                     _ptr(tbar)), 'oz_rhs_out_vjp')
    return yb, pbar, tbar


def rhs_jac_pattern(spec: SolveSpec, kind: int = KIND_TM):
    """oz_rhs_jac_pattern: boolean arrays (d, d) and (d, P) of the entries of df/dy, df/dp that may be nonzero."""
    import numpy as np
    pb = make_problem(spec, kind)
    d, P = spec.info['d'], spec.info['P']
    jp = (C.c_uint8 * (d * d))()
    jpp = (C.c_uint8 * max(d * P, 1))()
    nat.check(nat.lib.oz_rhs_jac_pattern(pb.rhs_id, jp, jpp), 'oz_rhs_jac_pattern')
    return (np.frombuffer(jp, dtype=np.uint8).reshape(d, d).astype(bool),
            np.frombuffer(jpp, dtype=np.uint8)[:d * P].reshape(d, P).astype(bool))


def rhs_jac_eval(spec: SolveSpec, y, p=None, th=None, want_f=True, want_p=True, compact=False, kind: int = KIND_TM,
                 stage_params=False):
    """oz_rhs_jac_eval on y (blocks, d, nodes), p (blocks, P, nodes) -> f (blocks, d, nodes), Jy (blocks, d, d, nodes),
    Jp (blocks, d, P, nodes) | None.  th (S,) or (S, nodes).  compact: Jy (blocks, nnz_y, nodes), Jp (blocks, nnz_p, nodes)
    hold only the entries rhs_jac_pattern() marks, in row-major order.  stage_params: blocks = n s and `p` is the RAW
    packed parameter array (n + 1, P, N) | (n + 1, P): the kernel interpolates it to the stage abscissae itself."""
    pb = make_problem(spec, kind)
    info, dev = spec.info, y.device
    blocks, _, nodes = y.shape
    y = y.contiguous()
    p = p.contiguous() if (p is not None and info['P']) else None
    th = th.contiguous() if (th is not None and info['S']) else None
    pb.th = _ptr(th)
    pb.th_shared = int(th is None or th.dim() == 1)
    d, P = info['d'], info['P']
    ny, npp = d * d, d * P
    if compact:
        jp, jpp = rhs_jac_pattern(spec, kind)
        ny, npp = int(jp.sum()), int(jpp.sum())
    f = torch.empty((blocks, d, nodes), dtype=spec.dtype, device=dev) if want_f else None
    Jy = torch.empty((blocks, ny, nodes) if compact else (blocks, d, d, nodes), dtype=spec.dtype, device=dev)
    Jp = torch.empty((blocks, npp, nodes) if compact else (blocks, d, P, nodes), dtype=spec.dtype, device=dev) if (P and want_p) else None
    nat.check(nat.lib.oz_rhs_jac_eval(C.byref(pb), nodes, blocks, int(compact) | (2 if stage_params else 0), _ptr(y), _ptr(p), _ptr(f),
                                      _ptr(Jy), _ptr(Jp)),
              'oz_rhs_jac_eval')
    return f, Jy, Jp


class _OutputsOnHistory(torch.autograd.Function):
    """Profile outputs at every grid point of a history and field outputs summed over time, on the tape
    (the reference's are csdl graph nodes, vector_based.py:33-49): backward is one oz_rhs_out_vjp launch."""

    @staticmethod
    def forward(ctx, spec, hist, p, th):
        nt, d, N = hist.shape
        info = spec.info
        y, pp, tt = _node_layout(spec, hist, p, th)
        _, prof, fld = rhs_eval(spec, y, pp, tt)
        ctx.spec, ctx.shape = spec, (nt, d, N)
        ctx.has = (p is not None and info['P'] > 0, th is not None and info['S'] > 0)
        ctx.save_for_backward(*[t for t in (y, pp, tt) if t is not None])
        e = lambda: torch.empty(0, dtype=spec.dtype, device=hist.device)
        prof = prof.reshape(info['nprof'], nt, N).permute(1, 0, 2).contiguous() if prof is not None else e()
        fld = fld.reshape(info['nfield'], nt, N).sum(dim=1) if fld is not None else e()
        return prof, fld

    @staticmethod
    def backward(ctx, g_prof, g_fld):
        spec, (nt, d, N) = ctx.spec, ctx.shape
        info = spec.info
        it = iter(ctx.saved_tensors)
        y = next(it)
        pp = next(it) if ctx.has[0] else None
        tt = next(it) if ctx.has[1] else None
        vp = vf = None
        if info['nprof'] and g_prof is not None and g_prof.numel():
            vp = g_prof.permute(1, 0, 2).reshape(info['nprof'], nt * N).contiguous()
        if info['nfield'] and g_fld is not None and g_fld.numel():      # the time sum broadcasts its cotangent
            vf = g_fld.unsqueeze(1).expand(info['nfield'], nt, N).reshape(info['nfield'], nt * N).contiguous()
        yb, pbar, tbar = rhs_out_vjp(spec, y, pp, tt, vp, vf)
        g_hist = yb.reshape(d, nt, N).permute(1, 0, 2).contiguous()
        g_p = g_th = None
        if pbar is not None and ctx.needs_input_grad[2]:
            g_p = pbar.reshape(info['P'], nt, N).permute(1, 0, 2)
            g_p = g_p.sum(dim=2) if spec.p_shared else g_p.contiguous()
        if tbar is not None and ctx.needs_input_grad[3]:
            g_th = tbar.reshape(info['S'], nt, N).sum(dim=1)
            g_th = g_th.sum(dim=1) if spec.th_shared else g_th
        return None, g_hist, g_p, g_th


def outputs_on_history(spec: SolveSpec, hist, p, th):
    """vector_based.py:33-49: profile outputs at every grid point of a solved history and field outputs
    summed over time, with the RAW parameters; differentiable w.r.t. the history, p and th."""
    prof, fld = _OutputsOnHistory.apply(spec, hist, p, th)
    return (prof if prof.numel() else None), (fld if fld.numel() else None)


class PipelinedTmSolve:
    """Solve + reverse sweep of a HOST-resident trajectory batch, cut into `slices` sub-batches that each run
    on their own stream: while one slice integrates, the next one's inputs cross PCIe and the previous one's
    gradients flow back (copies in both directions overlap each other and the kernels).  Trajectories are
    independent, so slicing changes nothing but the order of the sums of batch-shared gradients, which are
    added up over the slices at the end (fixed order)."""

    def __init__(self, spec: SolveSpec, slices=4, device='cuda', grads=('h', 'p'), cotangents=('final',),
                 upload=('h', 'p'), fractions=None):
        """`fractions`: relative sizes of the sub-batches (default: equal).  Decreasing sizes shorten the tail of the
        pipeline -- the last slice's kernels and download cannot overlap anything -- while the big early slices keep
        the kernels efficient; all uploads still queue back to back on the copy engine."""
        from .parallel import shard_bounds
        self.spec, self.upload = spec, tuple(upload)
        dev = torch.device(device)
        if fractions:
            tot, acc, cuts = float(sum(fractions)), 0.0, [0]
            for f in fractions[:-1]:
                acc += f
                cuts.append(min(spec.N, max(cuts[-1] + 1, int(round(spec.N * acc / tot / 64.0)) * 64)))
            cuts.append(spec.N)
            self.bounds = [(a, b) for a, b in zip(cuts[:-1], cuts[1:]) if b > a]
        else:
            slices = max(1, min(slices, spec.N))
            self.bounds = [shard_bounds(spec.N, i, slices) for i in range(slices)]
        self.plans = [TmPlan(replace(spec, N=hi - lo), device=dev, grads=grads, cotangents=cotangents)
                      for lo, hi in self.bounds]
        self.streams = [torch.cuda.Stream(device=dev) for _ in self.plans]
        self.shared = dict(y0=False, h=spec.h_shared, p=spec.p_shared, th=spec.th_shared)
        pin = lambda t: torch.empty(t.shape, dtype=t.dtype).pin_memory()
        self.host_in = [{k: pin(getattr(pl, k)) for k in ('y0', 'h', 'p', 'th') if getattr(pl, k) is not None}
                        for pl in self.plans]
        self.host_grad = [{k: pin(g) for k, g in pl.grad.items() if g is not None} for pl in self.plans]
        self.host_final = [pin(pl.final) for pl in self.plans]
        self._start, self._done = torch.cuda.Event(), [torch.cuda.Event() for _ in self.plans]
        self.host_shared = {k: pin(g) for k, g in self.plans[0].grad.items() if g is not None and self.shared[k]}
        self.reduce_shared = None      # e.g. torch.distributed.all_reduce under torchrun

    def set_host_inputs(self, **arrays):
        """arrays in the packed layout of the whole batch (torch CPU tensors); per-trajectory ones are split
        into contiguous per-slice pinned buffers once, inputs that are not re-uploaded go to the device now."""
        for (lo, hi), pl, hin in zip(self.bounds, self.plans, self.host_in):
            for k, a in arrays.items():
                if k not in hin:
                    continue
                hin[k].copy_(a if self.shared[k] else a[..., lo:hi])
                if k not in self.upload:
                    getattr(pl, k).copy_(hin[k])

    def capture(self):
        """Record the slice pipeline of run() once as a CUDA graph; later run() calls replay it.  A run is ~10
        enqueue calls per slice (copies, two solves, reducers, event bookkeeping), which at small sub-batches costs
        the host more time than the GPU needs to execute them; the replay is one launch.  Buffers are the plan's
        own, so the captured addresses stay valid.  Call after set_host_inputs()."""
        self.graph = None
        self.run()                                  # warm-up outside the capture (module loading, lazy init)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._enqueue_slices()
        self.graph = g
        return self

    def run(self):
        """Enqueue upload -> forward -> reverse -> download for every slice; returns immediately."""
        if getattr(self, 'graph', None) is not None:
            self.graph.replay()
        else:
            self._enqueue_slices()
        # gradients of batch-shared inputs: sum over the slices (fixed order), optional all-reduce over ranks, download
        for k in self.host_grad[0]:
            if self.shared[k]:
                total = torch.stack([pl.grad[k] for pl in self.plans]).sum(0)
                if self.reduce_shared is not None:
                    self.reduce_shared(total)
                self.host_shared[k].copy_(total, non_blocking=True)

    def _enqueue_slices(self):
        cur = torch.cuda.current_stream()
        self._start.record(cur)
        for pl, st, hin, hg, hf, done in zip(self.plans, self.streams, self.host_in, self.host_grad, self.host_final,
                                             self._done):
            st.wait_event(self._start)
            with torch.cuda.stream(st):
                for k in self.upload:
                    if k in hin:
                        getattr(pl, k).copy_(hin[k], non_blocking=True)
                pl.forward()
                pl.adjoint()
                hf.copy_(pl.final, non_blocking=True)
                for k, v in hg.items():
                    if not self.shared[k]:
                        v.copy_(pl.grad[k], non_blocking=True)
                done.record(st)
        for done in self._done:
            cur.wait_event(done)

    def bytes_per_run(self):
        w = 8 if self.spec.dtype == torch.float64 else 4
        h2d = sum(hin[k].numel() * w for hin in self.host_in for k in self.upload if k in hin)
        d2h = sum(v.numel() * w for hg in self.host_grad for k, v in hg.items() if not self.shared[k]) \
            + sum(v.numel() * w for v in self.host_shared.values()) + sum(f.numel() * w for f in self.host_final)
        return h2d, d2h

    def gather(self):
        """After a synchronisation: (final (d, N), grads) assembled on the host; shared gradients summed over slices."""
        final = torch.cat(self.host_final, dim=-1)
        grads = {}
        for k in self.host_grad[0]:
            grads[k] = self.host_shared[k] if self.shared[k] else torch.cat([hg[k] for hg in self.host_grad], dim=-1)
        return final, grads

    @property
    def launches(self):
        return sum(pl.launches_forward + pl.launches_adjoint for pl in self.plans)


class SolveStream:
    """A stream of INDEPENDENT host-resident batches of the same shape (an optimiser evaluating several designs, a service
    working through a queue): `depth` PipelinedTmSolve instances, each with its own device + pinned buffers and streams, used
    round-robin, so that solve k + 1 uploads while solve k computes and downloads and both copy engines stay busy.
    One solve alone is bound by the chain upload -> last sub-batch's kernels -> download; a stream of them by the link.

        q = SolveStream(spec, depth=3, slices=4, grads=('h', 'p'))
        q.set_host_inputs(slot, y0=..., h=..., p=...)        # for every slot, or slot=None for all
        q.begin();  slots = [q.submit() for _ in range(K)];  q.end()       # enqueue only; results after a synchronisation
        final, grads = q.gather(slot)                          # of the LAST solve submitted to that slot
    """

    def __init__(self, spec: SolveSpec, depth=3, device='cuda', **pipeline_args):
        self.pipes = [PipelinedTmSolve(spec, device=device, **pipeline_args) for _ in range(max(1, int(depth)))]
        self.streams = [torch.cuda.Stream(device=torch.device(device)) for _ in self.pipes]
        self._next = 0

    depth = property(lambda self: len(self.pipes))

    def set_host_inputs(self, slot=None, **arrays):
        for q, pipe in enumerate(self.pipes):
            if slot is None or q == slot:
                pipe.set_host_inputs(**arrays)

    def capture(self):
        for pipe in self.pipes:
            pipe.capture()
        return self

    def begin(self):
        """Order every slot after the work already enqueued on the current stream."""
        cur = torch.cuda.current_stream()
        for st in self.streams:
            st.wait_stream(cur)

    def submit(self):
        """Enqueue the next solve on the next slot (after the solve that used that slot before); returns the slot."""
        q = self._next
        self._next = (q + 1) % len(self.pipes)
        with torch.cuda.stream(self.streams[q]):
            self.pipes[q].run()
        return q

    def end(self):
        """Make the current stream wait for everything submitted."""
        cur = torch.cuda.current_stream()
        for st in self.streams:
            cur.wait_stream(st)

    def gather(self, slot):
        return self.pipes[slot].gather()


def measure_fma_peak(dtype=torch.float64, device='cuda', iters=20000, repeats=3):
    """Measured vector FMA peak of this GPU in TFLOP/s (oz_fma_probe timed with CUDA events, best of `repeats`)."""
    out = torch.zeros(1, dtype=dtype, device=device)
    code = nat.OZ_F64 if dtype == torch.float64 else nat.OZ_F32
    stream = torch.cuda.current_stream().cuda_stream
    nat.lib.oz_fma_probe(code, 200, out.data_ptr(), stream)           # warm-up
    best = 0.0
    for _ in range(repeats):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        flops = nat.lib.oz_fma_probe(code, iters, out.data_ptr(), stream)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, flops / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best
