#!/usr/bin/env python
"""Picard iteration with a large per-trajectory system: lunar ascent (7 states) x RK6 (7 stages), the instantiation
whose kernel keeps 2.7 kB of its state on the local-memory stack (profiles/r1l_resource_usage.txt).  Times the sweep
kernel and reports convergence for a few horizons.   python tools/time_picard_large.py [N]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from ozone_b200 import engine
from ozone_b200.glms import get_integration_method

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
nt = 101
dev = torch.device('cuda', 0)
for method in ('RK4', 'RK6'):
    for T in (0.05, 0.2, 0.5):
        wl = dict(kind='picard', rhs='lunar_ascent', method=method, N=N, nt=nt, T=T)
        case = bench.synth_inputs(wl, N, seed=0)
        t = lambda a: torch.tensor(np.ascontiguousarray(a), dtype=torch.float64, device=dev)
        spec = engine.SolveSpec(rhs_name='lunar_ascent', method=get_integration_method(method), N=N, n=nt - 1,
                                h_shared=True, p_shared=False, th_shared=True, state_blocks=(0, 7))
        y0, h, p = t(case['y0']), t(case['h']), t(case['p'])
        for _ in range(2):
            out = engine.picard_forward(spec, y0, h, p, None, tol=1e-10)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            out = engine.picard_forward(spec, y0, h, p, None, tol=1e-10)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        tm = engine.tm_forward(spec, y0, h, p, None)
        err = float((out['final'] - tm['final']).abs().max() / tm['final'].abs().max())
        print(f'{method} T={T}: {ms:8.2f} ms per solve, mean sweeps {float(out["iters"].double().mean()):.1f}, '
              f'not converged {int((out["status"] != 0).sum())} of {N}, |final - time marching| {err:.1e}, '
              f'{N * 7 * (nt - 1) / ms / 1e6:.2f} G state-steps/s', flush=True)
