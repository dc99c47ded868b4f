"""The peer-memory kernels (csrc/oz_peer.cu) with several ranks emulated on ONE GPU: their "peer" pointers are local
ones (peer.LocalArena), each emulated rank issues its kernels on its own stream so that they wait on each other the
way real ranks do.  The 2/4/8-GPU runs of the same code are driven by tools/check_pde_shard.py and bench.py."""
import numpy as np
import pytest
import torch

from ozone_b200 import peer

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('dtype', [torch.float64, torch.float32])
def test_peer_sum_over_emulated_ranks_is_the_rank_ordered_sum(dtype):
    world, rows = 4, 39
    arena = peer.LocalArena(1 << 16, 'cuda', world)
    sums, streams = [], [torch.cuda.Stream() for _ in range(world)]
    for r in range(world):
        v = arena.view(r)
        v._cursor = 0
        sums.append(peer.PeerSum(v, rows, dtype))
    g = torch.Generator(device='cuda').manual_seed(3)
    for call in range(3):                                  # three rounds: slots are reused with alternating parity
        local = [torch.randn(rows, dtype=dtype, device='cuda', generator=g) for _ in range(world)]
        out = [torch.full((rows,), float('nan'), dtype=dtype, device='cuda') for _ in range(world)]
        torch.cuda.synchronize()
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                sums[r](local[r], out[r], streams[r].cuda_stream)
        torch.cuda.synchronize()
        expect = local[0].clone()
        for r in range(1, world):
            expect = expect + local[r]
        for r in range(world):
            sums[r].check()
            assert torch.equal(out[r], expect), (call, r)


def test_halo_channel_over_emulated_ranks_and_timeout_flag():
    world, count = 3, 40
    arena = peer.LocalArena(1 << 16, 'cuda', world)
    chans = []
    for r in range(world):
        v = arena.view(r)
        v._cursor = 0
        chans.append(peer.HaloChannel(v, count, torch.float64))
    st = torch.cuda.current_stream().cuda_stream
    for epoch in range(3):
        owned = [torch.arange(count * 2, dtype=torch.float64, device='cuda') + 1000 * r + 10 * epoch for r in range(world)]
        halos = [(torch.zeros(count, dtype=torch.float64, device='cuda'), torch.zeros(count, dtype=torch.float64, device='cuda'))
                 for _ in range(world)]
        for phase in ('push', 'wait'):
            for r in range(world):
                chans[r].exchange(owned[r][:count], owned[r][count:], halos[r][0], halos[r][1], st, phase=phase)
        torch.cuda.synchronize()
        for r in range(world):
            chans[r].check()
            if r > 0:
                assert torch.equal(halos[r][0], owned[r - 1][count:])        # my left halo = left neighbour's right edge
            if r < world - 1:
                assert torch.equal(halos[r][1], owned[r + 1][:count])
    for c in chans:
        c.advance(st)
    torch.cuda.synchronize()
    assert all(c.epoch == 0 for c in chans)
    # and once more after the advance (what a replayed CUDA graph does)
    owned = [torch.full((2 * count,), float(r + 1), dtype=torch.float64, device='cuda') for r in range(world)]
    halos = [(torch.zeros(count, dtype=torch.float64, device='cuda'), torch.zeros(count, dtype=torch.float64, device='cuda'))
             for _ in range(world)]
    for phase in ('push', 'wait'):
        for r in range(world):
            chans[r].exchange(owned[r][:count], owned[r][count:], halos[r][0], halos[r][1], st, phase=phase)
    torch.cuda.synchronize()
    assert float(halos[1][0][0]) == 1.0 and float(halos[1][1][0]) == 3.0
