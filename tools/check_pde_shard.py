#!/usr/bin/env python
"""Run under torchrun on G GPUs: the spatially sharded diffusion-reaction solve (NCCL halo exchange)
against the undivided solve on rank 0.  Prints max relative differences and timings.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/check_pde_shard.py [nx n ni]
"""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ozone_b200 import engine, parallel
from ozone_b200.glms import get_integration_method

nx, n, ni = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (1_000_000, 1024, 32)
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
rng = np.random.default_rng(0)
dx = np.pi / nx
cfg = (dx, 16.0, 2.0, 2.0, float(round(0.3 * nx)), float(round(0.7 * nx)))
m = get_integration_method('RK4')
y0 = torch.full((nx,), 0.5, dtype=torch.float64, device=dev)
h = torch.full((n,), 0.4 * dx * dx, dtype=torch.float64, device=dev)
u = torch.tensor(rng.uniform(-0.6, 0.0, size=n + 1), dtype=torch.float64, device=dev)

times = []
for it in range(3):
    dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
    res = parallel.solve_grid_sharded(m, nx, n, ni, y0, h, u, cfg, device=dev)
    torch.cuda.synchronize(); dist.barrier(); times.append(time.perf_counter() - t0)
if rank == 0:
    spec = engine.SolveSpec(rhs_name='diffusion_reaction', method=m, N=1, n=n, h_shared=True, p_shared=True, th_shared=True,
                            num_inner=ni, store_history=False, d_override=nx, rhs_cfg=cfg)
    plan = engine.TmPlan(spec, device=dev, grads=('y0', 'h', 'p'), cotangents=('field',))
    plan.y0.copy_(y0.reshape(-1, 1)); plan.h.copy_(h); plan.p.copy_(u.reshape(-1, 1)); plan.field_bar.fill_(1.0)
    single = []
    for it in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        plan.forward(); plan.adjoint()
        torch.cuda.synchronize(); single.append(time.perf_counter() - t0)
    rel = lambda a, b: float((a - b).abs().max() / b.abs().max())
    lo, hi = res['own_range']
    print(f'world={world} nx={nx} n={n} ni={ni}  sharded {min(times) * 1e3:.2f} ms   single-GPU {min(single) * 1e3:.2f} ms')
    print('field   rel', rel(res['field'], plan.field))
    print('h_bar   rel', rel(res['h_bar'], plan.grad['h']))
    print('u_bar   rel', rel(res['u_bar'], plan.grad['p']))
    print('y0_bar  rel (owned slab)', rel(res['y0_bar_owned'], plan.grad['y0'][lo:hi, 0]))
    print('final   rel (owned slab)', rel(res['final_owned'], plan.final[lo:hi, 0]))
    print('cell-steps/s fwd+adj: sharded %.3g  single %.3g' % (nx * n / min(times), nx * n / min(single)))
dist.destroy_process_group()
