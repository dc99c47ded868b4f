"""Forward time per step of the fused grid kernel against the grid size (is a step latency- or throughput-bound?)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ozone_b200 import engine
from ozone_b200.glms import get_integration_method

m = get_integration_method('RK4')
n, ni = 256, 16
for nx in (125_000, 250_000, 500_000, 1_000_000, 2_000_000, 4_000_000):
    dx = np.pi / nx
    cfg = (dx, 16.0, 2.0, 2.0, float(round(0.3 * nx)), float(round(0.7 * nx)))
    y0 = torch.full((nx, 1), 0.5, dtype=torch.float64, device='cuda')
    h = torch.full((n,), 0.4 * dx * dx, dtype=torch.float64, device='cuda')
    u = torch.zeros(n + 1, 1, dtype=torch.float64, device='cuda')
    spec = engine.SolveSpec(rhs_name='diffusion_reaction', method=m, N=1, n=n, h_shared=True, p_shared=True, th_shared=True,
                            num_inner=ni, store_history=False, d_override=nx, rhs_cfg=cfg)
    for _ in range(2):
        engine.tm_forward(spec, y0, h, u, None, keep_history=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        engine.tm_forward(spec, y0, h, u, None, keep_history=False)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (3 * n)
    print(f'nx {nx:9d}: {us:7.2f} us/step  {nx / us * 1e-3:7.2f} Gcell-steps/s forward')
