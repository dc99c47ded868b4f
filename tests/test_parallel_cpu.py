"""Host-side logic of the multi-GPU layer on CPU: world_size-2 gloo processes (no GPU needed)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ozone_b200 import parallel


def test_shard_bounds_cover_the_batch_exactly():
    for total in (1, 7, 64, 65536, 1000003):
        for world in (1, 2, 3, 8):
            cuts = [parallel.shard_bounds(total, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1


def test_slab_partition_and_halo_copy_emulated():
    nx, world, halo = 103, 3, 8
    full = torch.arange(nx, dtype=torch.float64)
    slabs = [parallel.make_slab(nx, r, world, halo) for r in range(world)]
    exts = [full[s.ext_lo:s.ext_hi].clone() for s in slabs]
    for e, s in zip(exts, slabs):                 # wreck the halo zones, then restore them from the neighbours
        a, b = s.local_own
        e[:a] = -1
        e[b:] = -1
    pay = [parallel.halo_payloads(e, s) for e, s in zip(exts, slabs)]
    for r, (e, s) in enumerate(zip(exts, slabs)):
        parallel.set_halos(e, s, pay[r - 1][1] if r > 0 else None, pay[r + 1][0] if r < world - 1 else None)
        assert torch.equal(e, full[s.ext_lo:s.ext_hi])
    with pytest.raises(ValueError):
        parallel.make_slab(20, 0, 4, 8)           # slab narrower than its halo


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        # 1. trajectory batches: per-trajectory tensors are sliced, shared ones are not; shared grads are summed
        N = 11
        tensors = {'y0': torch.arange(2 * N, dtype=torch.float64).reshape(2, N), 'h': torch.ones(5, dtype=torch.float64)}
        local, (lo, hi) = parallel.shard_batch(tensors, {'h': True}, N, rank, world)
        assert local['y0'].shape == (2, hi - lo) and torch.equal(local['y0'], tensors['y0'][:, lo:hi])
        assert local['h'] is tensors['h']
        grads = {'h': torch.full((5,), float(rank + 1), dtype=torch.float64), 'y0': torch.full((2, hi - lo), 7.0, dtype=torch.float64)}
        parallel.allreduce_shared(grads, {'h': True})
        assert torch.equal(grads['h'], torch.full((5,), float(sum(range(1, world + 1))), dtype=torch.float64))
        assert torch.equal(grads['y0'], torch.full((2, hi - lo), 7.0, dtype=torch.float64))
        # 2. grid slabs: the halo exchange restores the neighbours' owned cells
        nx, halo = 64, 6
        full = torch.arange(nx, dtype=torch.float64) * 1.5
        slab = parallel.make_slab(nx, rank, world, halo)
        ext = full[slab.ext_lo:slab.ext_hi].clone()
        a, b = slab.local_own
        ext[:a] = -1.0
        ext[b:] = -1.0
        parallel.exchange_halos(ext, slab)
        assert torch.equal(ext, full[slab.ext_lo:slab.ext_hi])
        q.put((rank, 'ok'))
    except Exception as e:                         # pragma: no cover
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_sharding_and_halo_exchange():
    world, port = 2, _free_port()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, 'ok'), (1, 'ok')], res


def test_peer_arena_layout_without_a_gpu():
    """Address arithmetic of the peer-memory channels (ozone_b200/peer.py) on CPU buffers: every rank computes the same
    offsets, slots of different (parity, side) do not overlap, flags and slots stay inside the arena, 8-byte aligned."""
    from ozone_b200 import peer
    world, count = 4, 96
    arena = peer.LocalArena(4096 + 4 * count * 8, 'cpu', world)
    chans = []
    for r in range(world):
        v = arena.view(r)
        v._cursor = 0
        chans.append(peer.HaloChannel(v, count, torch.float64))
    assert len({(c.off_ctl, c.off_slots) for c in chans}) == 1            # same layout on every rank
    c = chans[1]
    spans = []
    for par in (0, 1):
        for side in (0, 1):
            a = c._slot(2, par, side) - arena.ptrs[2]
            spans.append((a, a + count * 8))
            assert a % 8 == 0 and 0 <= a and a + count * 8 <= arena.nbytes
    spans.sort()
    assert all(hi <= lo2 for (_, hi), (lo2, _) in zip(spans, spans[1:]))   # no overlap
    flags = [c._flag(2, s) - arena.ptrs[2] for s in (0, 1)]
    assert flags[0] != flags[1] and all(f % 8 == 0 and f + 8 <= c.off_slots for f in flags)
    with pytest.raises(ValueError):
        arena.view(0).region(1 << 30)                                      # the arena refuses to overflow
    v = arena.view(0)
    v._cursor = 0
    assert v.region(8 * world) == 0 and v.region(64) % 256 == 0
