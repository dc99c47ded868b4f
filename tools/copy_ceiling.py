#!/usr/bin/env python
"""Pinned host <-> device copy ceiling of this box, the denominator of the end-to-end leg (bench.py `e2e`).

    python tools/copy_ceiling.py [MB_h2d MB_d2h]          (under torchrun: every rank copies at once; rank 0 prints the per-rank figures)

One direction alone, then both directions at the same time on two streams (what the pipelined e2e leg does)."""
import os
import sys

import torch

mb_in, mb_out = (float(a) for a in sys.argv[1:3]) if len(sys.argv) >= 3 else (22.5, 22.5)
rank, world, local = (int(os.environ.get(k, d)) for k, d in (('RANK', 0), ('WORLD_SIZE', 1), ('LOCAL_RANK', 0)))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
hin = torch.empty(int(mb_in * 1e6), dtype=torch.uint8).pin_memory()
hout = torch.empty(int(mb_out * 1e6), dtype=torch.uint8).pin_memory()
din, dout = torch.empty_like(hin, device='cuda'), torch.empty_like(hout, device='cuda')
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
        s1.synchronize(); s2.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def h2d():
    with torch.cuda.stream(s1):
        din.copy_(hin, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        hout.copy_(dout, non_blocking=True)


def both():
    h2d(); d2h()


t_in, t_out, t_both = timed(h2d), timed(d2h), timed(both)
res = torch.tensor([mb_in / t_in, mb_out / t_out, (mb_in + mb_out) / t_both, t_both], dtype=torch.float64, device='cuda')
if world > 1:
    allr = [torch.zeros_like(res) for _ in range(world)]
    dist.all_gather(allr, res)
else:
    allr = [res]
if rank == 0:
    for r, v in enumerate(allr):
        a, b, c, t = v.tolist()
        print(f'rank {r}/{world}: H2D {a:.1f} GB/s  D2H {b:.1f} GB/s  both at once {c:.1f} GB/s combined ({t:.3f} ms for {mb_in} + {mb_out} MB)')
if world > 1:
    dist.destroy_process_group()
