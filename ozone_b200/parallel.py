"""Multi-GPU layer: one process per GPU, ``torch.distributed`` (NCCL over NVLink) for the plumbing.

The reference has no distributed code at all; the two sharding modes below are what
its "make the state bigger" idiom becomes on an 8-GPU box.

1. Trajectory batches (configs C2, C4, C5): the batch axis is cut into contiguous slices, one per
   rank.  The forward solve needs no communication.  In the reverse sweep only the gradients of
   batch-SHARED inputs (step sizes, shared parameters) are sums over trajectories, so they -- and
   nothing else -- are all-reduced (a few hundred scalars).

2. Spatial slabs of the diffusion-reaction grid (config C3): each rank owns a contiguous slab of
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

    ``forward_segment`` / ``adjoint_segment`` run one checkpoint segment on the extended slab through
    the ordinary C ABI in slab mode; between segments the caller exchanges halos (``halo_payloads`` /
    ``set_halos``, or ``exchange_halos`` under torch.distributed)."""

    def __init__(self, method, nx, n, num_inner, rank, world, y0_global, h, u, rhs_cfg, dtype=torch.float64,
                 device='cuda'):
        if n % num_inner:
            raise ValueError(f'Number of timesteps ({n}) must be divisible by the segment length {num_inner}')
        s = method.num_stages
        self.method, self.nx, self.n, self.ni, self.nseg = method, nx, n, num_inner, n // num_inner
        self.slab = make_slab(nx, rank, world, 2 * s * num_inner)
        self.dtype, self.device = dtype, torch.device(device)
        self.h = h.to(self.device, dtype).contiguous()
        self.u = u.to(self.device, dtype).reshape(n + 1, 1).contiguous()
        self.rhs_cfg = tuple(rhs_cfg)
        sl = self.slab
        self.ext = y0_global[sl.ext_lo:sl.ext_hi].to(self.device, dtype).reshape(-1, 1).contiguous()
        self.ckpts = [self.ext.clone()]
        self.field = torch.zeros(1, 1, dtype=dtype, device=self.device)
        self.h_bar = torch.zeros(n, dtype=dtype, device=self.device)
        self.u_bar = torch.zeros(n + 1, 1, dtype=dtype, device=self.device)
        self.lam = None

    def _spec(self, seg, adjoint):
        sl = self.slab
        a, b = sl.local_own
        flags = 0 if sl.rank == 0 else nat.GRID_NO_CONTROL_COST
        if not adjoint and seg > 0:
            flags |= nat.GRID_SKIP_IC_FIELD
        if adjoint and seg < self.nseg - 1:
            flags |= nat.GRID_SKIP_TERMINAL_FIELD
        return engine.SolveSpec(rhs_name='diffusion_reaction', method=self.method, N=1, n=self.ni, dtype=self.dtype,
                                h_shared=True, p_shared=True, th_shared=True, num_inner=self.ni, store_history=False,
                                d_override=sl.n_ext, rhs_cfg=self.rhs_cfg,
                                grid_slab=(sl.ext_lo, self.nx, a, b, flags))

    def forward_segment(self, seg):
        k0 = seg * self.ni
        out = engine.tm_forward(self._spec(seg, False), self.ext, self.h[k0:k0 + self.ni].contiguous(),
                                self.u[k0:k0 + self.ni + 1].contiguous(), None, keep_history=False)
        self.field += out['field']
        self.ext = out['final']          # owned cells exact; the halo zones must be refreshed before the next segment

    def end_forward_segment(self):
        self.ckpts.append(self.ext.clone())

    def start_adjoint(self, field_bar=1.0, final_bar_global=None):
        sl = self.slab
        self.field_bar = torch.full((1, 1), float(field_bar), dtype=self.dtype, device=self.device)
        self.lam = torch.zeros(sl.n_ext, 1, dtype=self.dtype, device=self.device)
        if final_bar_global is not None:
            self.lam += final_bar_global[sl.ext_lo:sl.ext_hi].to(self.device, self.dtype).reshape(-1, 1)

    def adjoint_segment(self, seg):
        k0 = seg * self.ni
        ck = torch.stack([self.ckpts[seg], self.ckpts[seg + 1]]).contiguous()
        g = engine.tm_adjoint(self._spec(seg, True), self.ckpts[seg], self.h[k0:k0 + self.ni].contiguous(),
                              self.u[k0:k0 + self.ni + 1].contiguous(), None, ckpt=ck, final_bar=self.lam,
                              field_bar=self.field_bar, need=('y0', 'h', 'p'))
        self.h_bar[k0:k0 + self.ni] = g['h']
        self.u_bar[k0:k0 + self.ni + 1] += g['p']
        self.lam = g['y0']               # owned cells exact; halo zones refreshed by the caller

    def owned(self, ext):
        a, b = self.slab.local_own
        return ext[a:b, 0]


def solve_grid_sharded(method, nx, n, num_inner, y0_global, h, u, rhs_cfg, field_bar=1.0, dtype=torch.float64,
                       device=None, group=None):
    """Distributed driver (call from every rank): forward + reverse sweep of the diffusion-reaction grid
    cut into slabs.  Returns dict(field, y0_bar_owned, own_range, h_bar, u_bar, final_owned)."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    device = device or torch.device('cuda', torch.cuda.current_device())
    sh = GridShard(method, nx, n, num_inner, rank, world, y0_global, h, u, rhs_cfg, dtype, device)
    for seg in range(sh.nseg):
        sh.forward_segment(seg)
        exchange_halos(sh.ext[:, 0], sh.slab, group)
        sh.end_forward_segment()
    final_owned = sh.owned(sh.ext).clone()
    sh.start_adjoint(field_bar)
    for seg in reversed(range(sh.nseg)):
        sh.adjoint_segment(seg)
        exchange_halos(sh.lam[:, 0], sh.slab, group)
    if world > 1:
        for t in (sh.field, sh.h_bar, sh.u_bar):
            dist.all_reduce(t, group=group)
    return dict(field=sh.field, y0_bar_owned=sh.owned(sh.lam).clone(), own_range=(sh.slab.own_lo, sh.slab.own_hi),
                h_bar=sh.h_bar, u_bar=sh.u_bar, final_owned=final_owned)


def solve_grid_emulated(method, nx, n, num_inner, world, y0_global, h, u, rhs_cfg, field_bar=1.0,
                        dtype=torch.float64, device='cuda'):
    """The same algorithm with all ``world`` ranks emulated in ONE process on one GPU (their kernels run
    one after another; the halo exchange is a copy).  Used to test the slab logic without N GPUs."""
    shards = [GridShard(method, nx, n, num_inner, r, world, y0_global, h, u, rhs_cfg, dtype, device) for r in range(world)]

    def swap(get):
        pay = [halo_payloads(get(s)[:, 0], s.slab) for s in shards]
        pay = [(None if l is None else l.clone(), None if r is None else r.clone()) for l, r in pay]
        for r, s in enumerate(shards):
            set_halos(get(s)[:, 0], s.slab, pay[r - 1][1] if r > 0 else None, pay[r + 1][0] if r < world - 1 else None)

    for seg in range(shards[0].nseg):
        for s in shards:
            s.forward_segment(seg)
        swap(lambda s: s.ext)
        for s in shards:
            s.end_forward_segment()
    final = torch.cat([s.owned(s.ext) for s in shards])
    for s in shards:
        s.start_adjoint(field_bar)
    for seg in reversed(range(shards[0].nseg)):
        for s in shards:
            s.adjoint_segment(seg)
        swap(lambda s: s.lam)
    return dict(field=sum(s.field for s in shards), y0_bar=torch.cat([s.owned(s.lam) for s in shards]),
                h_bar=sum(s.h_bar for s in shards), u_bar=sum(s.u_bar for s in shards), final=final)
