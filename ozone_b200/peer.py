"""Peer-mapped memory for the multi-GPU paths (one process per GPU on one NVLink / NVSwitch box).

``PeerArena`` allocates one buffer per rank in torch's symmetric memory, so that every rank holds a device
pointer to every other rank's buffer, and hands out regions of it (same offsets on all ranks).  The kernels
behind ``oz_peer_*`` (csrc/oz_peer.cu) then move halos and tiny gradient sums with plain loads / stores and
epoch flags instead of one NCCL call each; torch.distributed stays the plumbing (rendezvous, barriers, the
larger all-reduces).  ``LocalArena`` gives the same interface inside ONE process -- several emulated ranks on
one GPU, the "peer" pointers being local ones -- which is how the slab logic is tested without N GPUs.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _native as nat


def _align(x, a=256):
    return (x + a - 1) // a * a


class _ArenaBase:
    """Bump allocator over a byte buffer that exists once per rank with the same layout everywhere."""

    def __init__(self):
        self._cursor = 0

    def region(self, nbytes):
        """-> byte offset of a fresh region (call in the same order with the same sizes on every rank)."""
        off = self._cursor
        self._cursor = _align(off + nbytes)
        if self._cursor > self.nbytes:
            raise ValueError(f'peer arena of {self.nbytes} bytes is too small')
        return off

    def addr(self, rank, off):
        return self.ptrs[rank] + off

    def local(self, off, count, dtype):
        """Tensor view of a region of this rank's own buffer."""
        w = torch.empty(0, dtype=dtype).element_size()
        return self.buf[off:off + count * w].view(dtype)


class PeerArena(_ArenaBase):
    def __init__(self, nbytes, device, group=None):
        super().__init__()
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        group = group or dist.group.WORLD
        self.nbytes = _align(nbytes)
        self.buf = symm.empty(self.nbytes, dtype=torch.uint8, device=device)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        self.handle = symm.rendezvous(self.buf, group)
        self.rank, self.world = self.handle.rank, self.handle.world_size
        self.ptrs = [int(p) for p in self.handle.buffer_ptrs]
        self.handle.barrier()                    # every arena is zeroed before anybody signals into it

    def barrier(self):
        torch.cuda.synchronize()
        self.handle.barrier()


class LocalArena(_ArenaBase):
    """`world` arenas in one process on one device (emulated ranks)."""

    def __init__(self, nbytes, device, world):
        super().__init__()
        self.nbytes = _align(nbytes)
        self.bufs = [torch.zeros(self.nbytes, dtype=torch.uint8, device=device) for _ in range(world)]
        self.ptrs = [b.data_ptr() for b in self.bufs]
        self.world = world

    def view(self, rank):
        v = _ArenaBase.__new__(LocalArena)
        v.__dict__.update(self.__dict__)
        v.rank, v.buf = rank, self.bufs[rank]
        return v

    def barrier(self):
        torch.cuda.synchronize()


class HaloChannel:
    """Neighbour exchange of `count` values per side between consecutive ranks, double-buffered by epoch parity.

    Arena layout per rank: flags [from_left, from_right] (u64), epoch_base (u64), err (i32), then slots
    [parity][side][count] where side 0 receives from the left neighbour and side 1 from the right one."""

    def __init__(self, arena, count, dtype):
        self.arena, self.count, self.dtype = arena, int(count), dtype
        self.w = torch.empty(0, dtype=dtype).element_size()
        self.code = nat.OZ_F64 if dtype == torch.float64 else nat.OZ_F32
        self.off_ctl = arena.region(64)
        self.off_slots = arena.region(4 * self.count * self.w)
        self.rank, self.world = arena.rank, arena.world
        self.epoch = 0                           # epochs issued since the last advance (relative to *epoch_base)
        self.fused = isinstance(arena, PeerArena)   # one launch per exchange when every rank is its own process / GPU
        self.err = arena.local(self.off_ctl + 24, 1, torch.int32)

    def _slot(self, rank, parity, side):
        return self.arena.addr(rank, self.off_slots + ((parity * 2 + side) * self.count) * self.w)

    def _flag(self, rank, side):
        return self.arena.addr(rank, self.off_ctl + 8 * side)

    def exchange(self, left_src, right_src, left_halo, right_halo, stream, phase='both'):
        """left_src / right_src: this rank's owned edge cells (what the left / right neighbour keeps as halo);
        left_halo / right_halo: where the neighbours' cells go.  All tensors of `count` values; entries for an
        absent neighbour are ignored.  phase: 'push', 'wait' or 'both' (emulated ranks push first, then wait)."""
        r, last = self.rank, self.world - 1
        e = self.epoch + 1
        par = e & 1
        base = self.arena.addr(r, self.off_ctl + 16)
        has_l, has_r = r > 0, r < last
        if phase == 'both' and self.fused:
            nat.check(nat.lib.oz_peer_halo_exchange(
                self.code, left_src.data_ptr() if has_l else None, right_src.data_ptr() if has_r else None, self.count,
                self._slot(r - 1, par, 1) if has_l else None, self._slot(r + 1, par, 0) if has_r else None,
                self._flag(r - 1, 1) if has_l else None, self._flag(r + 1, 0) if has_r else None,
                self._slot(r, par, 0) if has_l else None, self._slot(r, par, 1) if has_r else None,
                left_halo.data_ptr() if has_l else None, right_halo.data_ptr() if has_r else None,
                self._flag(r, 0) if has_l else None, self._flag(r, 1) if has_r else None,
                base, e, self.err.data_ptr(), stream), 'oz_peer_halo_exchange')
            self.epoch = e
            return
        if phase in ('push', 'both'):
            nat.check(nat.lib.oz_peer_halo_push(
                self.code, left_src.data_ptr() if has_l else None, right_src.data_ptr() if has_r else None, self.count,
                self._slot(r - 1, par, 1) if has_l else None,        # I am my left neighbour's RIGHT source
                self._slot(r + 1, par, 0) if has_r else None,
                self._flag(r - 1, 1) if has_l else None, self._flag(r + 1, 0) if has_r else None,
                base, e, stream), 'oz_peer_halo_push')
        if phase in ('wait', 'both'):
            nat.check(nat.lib.oz_peer_halo_wait(
                self.code, self._slot(r, par, 0) if has_l else None, self._slot(r, par, 1) if has_r else None, self.count,
                left_halo.data_ptr() if has_l else None, right_halo.data_ptr() if has_r else None,
                self._flag(r, 0) if has_l else None, self._flag(r, 1) if has_r else None,
                base, e, self.err.data_ptr(), stream), 'oz_peer_halo_wait')
            self.epoch = e

    def advance(self, stream):
        """End of a run: move the epoch base so that the same (captured) sequence can be issued again."""
        if self.epoch:
            nat.check(nat.lib.oz_peer_epoch_advance(self.arena.addr(self.rank, self.off_ctl + 16), self.epoch, stream),
                      'oz_peer_epoch_advance')
        self.epoch = 0

    def check(self):
        if int(self.err.item()) != 0:
            raise RuntimeError('halo exchange timed out: a neighbour rank never signalled (did it fail?)')


class PeerSum:
    """Sum of `rows` values over all ranks, in rank order, as one small kernel (no NCCL call)."""

    def __init__(self, arena, rows, dtype):
        self.arena, self.rows, self.dtype = arena, int(rows), dtype
        self.w = torch.empty(0, dtype=dtype).element_size()
        self.code = nat.OZ_F64 if dtype == torch.float64 else nat.OZ_F32
        world = arena.world
        self.off_flags = arena.region(8 * world)
        self.off_ctl = arena.region(64)          # epoch_base (u64), err (i32)
        self.off_slots = arena.region(2 * world * self.rows * self.w)
        dev = arena.buf.device
        self.slot_ptrs = torch.tensor([arena.addr(r, self.off_slots) for r in range(world)], dtype=torch.int64, device=dev)
        self.flag_ptrs = torch.tensor([arena.addr(r, self.off_flags) for r in range(world)], dtype=torch.int64, device=dev)
        self.err = arena.local(self.off_ctl + 8, 1, torch.int32)
        self.epoch = 0

    def __call__(self, local, out, stream):
        """out[j] = sum over ranks of local[j]; `local` and `out` may be the same tensor."""
        self.epoch += 1
        nat.check(nat.lib.oz_peer_sum(self.code, local.data_ptr(), self.rows, self.arena.rank, self.arena.world,
                                      self.slot_ptrs.data_ptr(), self.flag_ptrs.data_ptr(),
                                      self.arena.addr(self.arena.rank, self.off_ctl), self.epoch, out.data_ptr(),
                                      self.err.data_ptr(), stream), 'oz_peer_sum')

    def check(self):
        if int(self.err.item()) != 0:
            raise RuntimeError('peer sum timed out: a rank never signalled (did it fail?)')
