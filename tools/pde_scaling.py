#!/usr/bin/env python
"""Per-step time of the fused grid step as a function of the grid size (is it throughput- or latency-bound?)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ozone_b200 import engine
from ozone_b200.glms import get_integration_method
m = get_integration_method('RK4')
n = 256
sizes = [int(a) for a in sys.argv[1].split(',')] if len(sys.argv) > 1 else (125_000, 250_000, 500_000, 1_000_000, 2_000_000, 4_000_000, 8_000_000)
print('fuse', os.environ.get('OZ_PDE_FUSE_F'), os.environ.get('OZ_PDE_FUSE_A'))
for dtype in (torch.float64, torch.float32):
    for nx in sizes:
        cfg = (np.pi / nx, 16.0, 2.0, 2.0, float(round(0.3 * nx)), float(round(0.7 * nx)))
        spec = engine.SolveSpec(rhs_name='diffusion_reaction', method=m, N=1, n=n, dtype=dtype, h_shared=True, p_shared=True,
                                th_shared=True, num_inner=32, store_history=False, d_override=nx, rhs_cfg=cfg)
        plan = engine.TmPlan(spec, grads=('y0', 'h', 'p'), cotangents=('field',))
        plan.y0.fill_(0.5); plan.h.fill_(0.4 * (np.pi / nx) ** 2); plan.p.fill_(-0.3); plan.field_bar.fill_(1.0)
        for _ in range(2):
            plan.forward(); plan.adjoint()
        torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record(); plan.forward(); e[1].record(); plan.adjoint(); e[2].record(); torch.cuda.synchronize()
        f, a = e[0].elapsed_time(e[1]) * 1e3 / n, e[1].elapsed_time(e[2]) * 1e3 / n
        print(f'{str(dtype)[6:]:8s} nx={nx:9d}  forward {f:7.2f} us/step ({nx / f / 1e3:6.1f} Gcell-steps/s)   reverse+replay {a:7.2f} us/step')
