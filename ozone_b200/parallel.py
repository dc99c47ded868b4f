"""Multi-GPU layer: one process per GPU, ``torch.distributed`` (NCCL over NVLink) for the plumbing.

The reference has no distributed code at all; the two sharding modes below are what
its "make the state bigger" idiom becomes on an 8-GPU box.

1. Trajectory batches (configs C2, C4, C5): the batch axis is cut into contiguous slices, one per
   rank.  The forward solve needs no communication.  In the reverse sweep only the gradients of
   batch-SHARED inputs (step sizes, shared parameters) are sums over trajectories, so they -- and
   nothing else -- are all-reduced (a few hundred scalars).

2. Spatial slabs of a grid right-hand side (config C3: the diffusion-reaction system; a traced grid callback works the same): each rank owns a contiguous slab of
   cells and carries a halo on either side.  An RK step with s stages moves information s cells
   (forward) / 2 s cells (reverse step, which re-derives the stages and then runs the transposed
   stage chain), so instead of exchanging one cell per stage the ranks exchange a halo of
   ``2 s n_inner`` cells once per checkpoint segment and compute redundantly on it: a slab treated
   as a grid with zero ghosts is exact on its owned cells until the error entering at its ends has
   travelled through the halo.  Grid-wide sums (J_space, u_bar, h_bar) are formed over owned cells
   and all-reduced once at the end.
"""
from __future__ import annotations

from dataclasses import dataclass, replace
from typing import Optional

import torch
import torch.distributed as dist

from . import engine
from . import _native as nat


# ------------------------------------------------------------------------------------------------
def shard_bounds(total: int, rank: int, world: int):
    """Contiguous, balanced partition of range(total): the first ``total % world`` ranks get one more."""
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_batch(tensors: dict, shared: dict, N: int, rank: int, world: int):
    """Slice the trailing (trajectory) axis of every per-trajectory tensor; shared tensors pass through."""
    lo, hi = shard_bounds(N, rank, world)
    out = {}
    for k, t in tensors.items():
        out[k] = t if (t is None or shared.get(k, False)) else t[..., lo:hi].contiguous()
    return out, (lo, hi)


def allreduce_shared(grads: dict, shared: dict, group=None):
    """Sum over ranks the gradients of batch-shared inputs (in place); per-trajectory ones stay local."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return grads
    for k, g in grads.items():
        if g is not None and shared.get(k, False):
            dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group)
    return grads


# ------------------------------------------------------------------------------------------------
@dataclass
class Slab:
    """Cells [own_lo, own_hi) of a grid of nx cells belong to this rank; it stores [ext_lo, ext_hi)."""
    nx: int
    rank: int
    world: int
    halo: int
    own_lo: int
    own_hi: int
    ext_lo: int
    ext_hi: int

    @property
    def n_ext(self):
        return self.ext_hi - self.ext_lo

    @property
    def local_own(self):
        return self.own_lo - self.ext_lo, self.own_hi - self.ext_lo


def make_slab(nx: int, rank: int, world: int, halo: int) -> Slab:
    lo, hi = shard_bounds(nx, rank, world)
    if world > 1 and (hi - lo) < halo:
        raise ValueError(f'slab of {hi - lo} cells is narrower than the halo ({halo}); use fewer ranks or shorter segments')
    return Slab(nx, rank, world, halo, lo, hi, max(0, lo - halo), min(nx, hi + halo))


def halo_payloads(ext: torch.Tensor, slab: Slab):
    """(to_left, to_right): this rank's owned edge cells that its neighbours keep as halo."""
    a, b = slab.local_own
    left = ext[a:a + slab.halo] if slab.rank > 0 else None
    right = ext[b - slab.halo:b] if slab.rank < slab.world - 1 else None
    return left, right


def set_halos(ext: torch.Tensor, slab: Slab, from_left: Optional[torch.Tensor], from_right: Optional[torch.Tensor]):
    a, b = slab.local_own
    if from_left is not None:
        ext[a - slab.halo:a] = from_left
    if from_right is not None:
        ext[b:b + slab.halo] = from_right


def exchange_halos(ext: torch.Tensor, slab: Slab, group=None):
    """Neighbour exchange of the halo zones (batched isend/irecv: NCCL on GPUs, gloo on CPU tensors)."""
    if slab.world == 1:
        return
    to_left, to_right = halo_payloads(ext, slab)
    ops, from_left, from_right = [], None, None
    if slab.rank > 0:
        from_left = torch.empty(slab.halo, dtype=ext.dtype, device=ext.device)
        ops += [dist.P2POp(dist.isend, to_left.contiguous(), slab.rank - 1, group),
                dist.P2POp(dist.irecv, from_left, slab.rank - 1, group)]
    if slab.rank < slab.world - 1:
        from_right = torch.empty(slab.halo, dtype=ext.dtype, device=ext.device)
        ops += [dist.P2POp(dist.isend, to_right.contiguous(), slab.rank + 1, group),
                dist.P2POp(dist.irecv, from_right, slab.rank + 1, group)]
    for req in dist.batch_isend_irecv(ops):
        req.wait()
    set_halos(ext, slab, from_left, from_right)


class GridShard:
    """One rank's part of a checkpointed diffusion-reaction solve (TimeMarchingCheckpoints, explicit method).

    Everything is allocated once: the checkpoints of the extended slab ``ckpts [nseg + 1][n_ext]``, the multiplier
    ping-pong, per-segment gradient rows, one workspace, and one prebuilt C-ABI descriptor per (segment, direction),
    so a solve only enqueues kernels -- the whole forward + reverse sweep can be captured in a CUDA graph when the
    halo exchange is the peer-memory one (``channel``, peer.HaloChannel); with torch.distributed P2P (``group``, also
    the gloo path on CPU tensors) the exchange is a batched isend / irecv per segment."""

    def __init__(self, method, nx, n, num_inner, rank, world, y0_global, h, u, rhs_cfg, dtype=torch.float64,
                 device='cuda', channel=None, traced=None):
        """traced: a trace_grid.TracedGridRHS instead of the built-in model (rhs_cfg is then unused); its coefficient
        fields are indexed by the GLOBAL cell, so every rank keeps the whole table."""
        import ctypes as C
        if n % num_inner:
            raise ValueError(f'Number of timesteps ({n}) must be divisible by the segment length {num_inner}')
        s = method.num_stages
        self.method, self.nx, self.n, self.ni, self.nseg = method, nx, n, num_inner, n // num_inner
        self.slab = sl = make_slab(nx, rank, world, 2 * s * num_inner)
        self.dtype, self.device = dtype, torch.device(device)
        self.h = h.to(self.device, dtype).contiguous()
        self.u = u.to(self.device, dtype).reshape(n + 1, -1).contiguous()       # (n + 1, P): one row of controls per time point
        P = self.u.shape[1]
        self.rhs_cfg = tuple(rhs_cfg or ())
        self.channel = channel
        self.traced, self.coef = traced, None
        if traced is not None and traced.tracer.coefs:
            self.coef = torch.from_numpy(traced.tracer.coefficient_table().reshape(-1)).to(self.device, dtype).contiguous()
        ne, ni, nseg = sl.n_ext, num_inner, self.nseg
        new = lambda *shape: torch.zeros(shape, dtype=dtype, device=self.device)
        self.ckpts = new(nseg + 1, ne, 1)
        self.ckpts[0] = y0_global[sl.ext_lo:sl.ext_hi].to(self.device, dtype).reshape(-1, 1)
        self.lam = [new(ne, 1), new(ne, 1)]
        self.field_seg, self.field_bar = new(nseg, 1, 1), new(1, 1)
        self.h_bar, self.u_seg = new(n), new(nseg, ni + 1, P)
        self._final_bar = None
        # descriptors
        self.pb_f, self.pb_a = [], []
        ws_bytes = 8
        for seg in range(nseg):
            k0 = seg * ni
            for adjoint in (False, True):
                pb = engine.make_problem(self._spec(seg, adjoint))
                pb.y0, pb.h, pb.p = self.ckpts[seg].data_ptr(), self.h[k0:].data_ptr(), self.u[k0:].data_ptr()
                if self.coef is not None:
                    pb.th, pb.th_shared = self.coef.data_ptr(), 1
                if adjoint:
                    pb.ckpt = self.ckpts[seg].data_ptr()                 # rows seg, seg + 1: the segment's two checkpoints
                    pb.final_bar, pb.y0_bar = None, None                 # set per call (ping-pong)
                    pb.field_bar = self.field_bar.data_ptr()
                    pb.h_bar, pb.p_bar = self.h_bar[k0:].data_ptr(), self.u_seg[seg].data_ptr()
                    need = nat.lib.oz_tm_adjoint_workspace_bytes(C.byref(pb))
                else:
                    pb.yfinal, pb.field = self.ckpts[seg + 1].data_ptr(), self.field_seg[seg].data_ptr()
                    need = nat.lib.oz_tm_forward_workspace_bytes(C.byref(pb))
                if need == C.c_size_t(-1).value:
                    raise nat.OzError(nat.lib.oz_last_error().decode())
                ws_bytes = max(ws_bytes, need)
                (self.pb_a if adjoint else self.pb_f).append(pb)
        self.ws = torch.empty(ws_bytes, dtype=torch.uint8, device=self.device)
        for pb in self.pb_f + self.pb_a:
            pb.ws, pb.ws_bytes = self.ws.data_ptr(), ws_bytes
        self.cur = 0
        self.launches_per_solve = 0

    def _spec(self, seg, adjoint):
        sl = self.slab
        a, b = sl.local_own
        flags = 0 if sl.rank == 0 else nat.GRID_NO_CONTROL_COST
        if not adjoint and seg > 0:
            flags |= nat.GRID_SKIP_IC_FIELD
        if adjoint and seg < self.nseg - 1:
            flags |= nat.GRID_SKIP_TERMINAL_FIELD
        return engine.SolveSpec(rhs_name='traced' if self.traced is not None else 'diffusion_reaction', traced=self.traced,
                                method=self.method, N=1, n=self.ni, dtype=self.dtype,
                                h_shared=True, p_shared=True, th_shared=True, num_inner=self.ni, store_history=False,
                                d_override=sl.n_ext, rhs_cfg=() if self.traced is not None else self.rhs_cfg,
                                grid_slab=(sl.ext_lo, self.nx, a, b, flags))

    # ---- halo exchange of an extended-slab vector (after a segment: owned cells exact, halo zones stale)
    def halo_views(self, ext):
        """(left_src, right_src, left_halo, right_halo) views of `ext` (n_ext, 1): what the neighbours need / refresh."""
        sl = self.slab
        a, b = sl.local_own
        v, hw = ext[:, 0], sl.halo
        return (v[a:a + hw] if sl.rank > 0 else None, v[b - hw:b] if sl.rank < sl.world - 1 else None,
                v[a - hw:a] if sl.rank > 0 else None, v[b:b + hw] if sl.rank < sl.world - 1 else None)

    def exchange(self, ext, group=None, phase='both'):
        sl = self.slab
        if sl.world == 1:
            return
        if self.channel is not None:
            ls, rs, lh, rh = self.halo_views(ext)
            self.channel.exchange(ls, rs, lh, rh, torch.cuda.current_stream().cuda_stream, phase=phase)
        elif phase in ('wait', 'both'):
            exchange_halos(ext[:, 0], sl, group)

    # ---- one segment each way (kernels only)
    def forward_segment(self, seg):
        import ctypes as C
        pb = self.pb_f[seg]
        pb.stream = torch.cuda.current_stream().cuda_stream
        nat.check(nat.lib.oz_tm_forward(C.byref(pb)), 'oz_tm_forward')

    def start_adjoint(self, field_bar=1.0, final_bar_global=None):
        sl = self.slab
        self.field_bar.fill_(float(field_bar))
        self.cur = 0
        self.lam[0].zero_()
        if final_bar_global is not None:
            self.lam[0] += final_bar_global[sl.ext_lo:sl.ext_hi].to(self.device, self.dtype).reshape(-1, 1)

    def adjoint_segment(self, seg):
        import ctypes as C
        pb = self.pb_a[seg]
        pb.stream = torch.cuda.current_stream().cuda_stream
        pb.final_bar, pb.y0_bar = self.lam[self.cur].data_ptr(), self.lam[self.cur ^ 1].data_ptr()
        nat.check(nat.lib.oz_tm_adjoint(C.byref(pb)), 'oz_tm_adjoint')
        self.cur ^= 1

    @property
    def lam_cur(self):
        return self.lam[self.cur]

    @property
    def ext(self):                       # state of the extended slab after the last forward segment
        return self.ckpts[self.nseg]

    def owned(self, ext):
        a, b = self.slab.local_own
        return ext[a:b, 0]

    # ---- whole solve
    def run(self, group=None, field_bar=1.0):
        """Forward sweep, then reverse sweep; kernels (and peer-memory halo kernels) only when `channel` is set."""
        for seg in range(self.nseg):
            self.forward_segment(seg)
            self.exchange(self.ckpts[seg + 1], group)
        self.start_adjoint(field_bar)
        for seg in reversed(range(self.nseg)):
            self.adjoint_segment(seg)
            self.exchange(self.lam_cur, group)
        if self.channel is not None:
            self.channel.advance(torch.cuda.current_stream().cuda_stream)

    def results(self):
        """Local (not yet all-reduced) field, h_bar, u_bar and the owned parts of the final state and of y0_bar."""
        n, ni, nseg = self.n, self.ni, self.nseg
        P = self.u.shape[1]
        u_bar = torch.zeros(n + 1, P, dtype=self.dtype, device=self.device)
        u_bar[:n].view(nseg, ni, P).add_(self.u_seg[:, :ni])
        u_bar[ni::ni].add_(self.u_seg[:, ni])            # the last grid point of a segment is the first of the next
        return dict(field=self.field_seg.sum(dim=0), h_bar=self.h_bar, u_bar=u_bar,
                    final_owned=self.owned(self.ext), y0_bar_owned=self.owned(self.lam_cur))


def make_channel(method, nx, num_inner, dtype, device, group=None):
    """Peer-memory halo channel for the spatial shard (CUDA tensors, >1 rank, symmetric memory available), else None."""
    from . import peer
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return None
    if torch.device(device).type != 'cuda':
        return None
    halo = 2 * method.num_stages * num_inner
    w = 8 if dtype == torch.float64 else 4
    arena = peer.PeerArena(4096 + 4 * halo * w, device, group)
    return peer.HaloChannel(arena, halo, dtype)


def solve_grid_sharded(method, nx, n, num_inner, y0_global, h, u, rhs_cfg, field_bar=1.0, dtype=torch.float64,
                       device=None, group=None, peer_halos=True, traced=None):
    """Distributed driver (call from every rank): forward + reverse sweep of the diffusion-reaction grid
    cut into slabs.  Returns dict(field, y0_bar_owned, own_range, h_bar, u_bar, final_owned)."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    device = device or torch.device('cuda', torch.cuda.current_device())
    channel = make_channel(method, nx, num_inner, dtype, device, group) if peer_halos else None
    sh = GridShard(method, nx, n, num_inner, rank, world, y0_global, h, u, rhs_cfg, dtype, device, channel=channel, traced=traced)
    sh.run(group, field_bar)
    res = sh.results()
    if channel is not None:
        torch.cuda.synchronize()
        channel.check()
    if world > 1:
        for t in (res['field'], res['h_bar'], res['u_bar']):
            dist.all_reduce(t, group=group)
    return dict(field=res['field'], y0_bar_owned=res['y0_bar_owned'].clone(), own_range=(sh.slab.own_lo, sh.slab.own_hi),
                h_bar=res['h_bar'], u_bar=res['u_bar'], final_owned=res['final_owned'].clone(), shard=sh)


def solve_grid_emulated(method, nx, n, num_inner, world, y0_global, h, u, rhs_cfg, field_bar=1.0,
                        dtype=torch.float64, device='cuda', peer_halos=True, traced=None):
    """The same algorithm with all ``world`` ranks emulated in ONE process on one GPU (their kernels run
    one after another).  With peer_halos the exchange goes through the oz_peer_* kernels over a LocalArena --
    every rank pushes, then every rank waits -- else it is a tensor copy.  Tests the slab logic without N GPUs."""
    from . import peer
    channels = [None] * world
    if peer_halos and world > 1:
        halo = 2 * method.num_stages * num_inner
        arena = peer.LocalArena(4096 + 4 * halo * (8 if dtype == torch.float64 else 4), device, world)
        for r in range(world):
            view = arena.view(r)
            view._cursor = 0
            channels[r] = peer.HaloChannel(view, halo, dtype)
    shards = [GridShard(method, nx, n, num_inner, r, world, y0_global, h, u, rhs_cfg, dtype, device, channel=channels[r],
                        traced=traced)
              for r in range(world)]

    def swap(get):
        if peer_halos and world > 1:
            for ph in ('push', 'wait'):
                for s in shards:
                    s.exchange(get(s), phase=ph)
            return
        pay = [halo_payloads(get(s)[:, 0], s.slab) for s in shards]
        pay = [(None if l is None else l.clone(), None if r is None else r.clone()) for l, r in pay]
        for r, s in enumerate(shards):
            set_halos(get(s)[:, 0], s.slab, pay[r - 1][1] if r > 0 else None, pay[r + 1][0] if r < world - 1 else None)

    nseg = shards[0].nseg
    for seg in range(nseg):
        for s in shards:
            s.forward_segment(seg)
        swap(lambda s: s.ckpts[seg + 1])
    for s in shards:
        s.start_adjoint(field_bar)
    for seg in reversed(range(nseg)):
        for s in shards:
            s.adjoint_segment(seg)
        swap(lambda s: s.lam_cur)
    res = [s.results() for s in shards]
    if peer_halos and world > 1:
        torch.cuda.synchronize()
        for c in channels:
            c.check()
    return dict(field=sum(r['field'] for r in res), y0_bar=torch.cat([r['y0_bar_owned'] for r in res]),
                h_bar=sum(r['h_bar'] for r in res), u_bar=sum(r['u_bar'] for r in res),
                final=torch.cat([r['final_owned'] for r in res]))
