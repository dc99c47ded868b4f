#!/usr/bin/env python
"""Launch each kernel family a few times at its bench size so that `ncu --set full` sees a short launch list.

    python tools/ncu_targets.py vdp|lv|lv32|pde|pde32|pdetr|colloc|coljac|picard

(The bench itself issues thousands of launches for the grid workload; profiling it directly is slow.)
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from ozone_b200 import engine

which = sys.argv[1]
name = {'vdp': 'van_der_pol_gl6_tm_adjoint_N65536_nt40', 'lv': 'lotka_volterra_rk4_tm_adjoint_N1M_nt101',
        'lv32': 'lotka_volterra_rk4_tm_adjoint_N1M_nt101_fp32', 'pde': 'diffusion_reaction_rk4_checkpoints_nx1M_n1024',
        'pde32': 'diffusion_reaction_rk4_checkpoints_nx1M_n1024_fp32',
        'pdetr': 'diffusion_reaction_rk4_checkpoints_nx1M_n1024_traced',     # the same PDE as a traced Python callback
        'colloc': 'lunar_ascent_rk4_collocation_N16384_nt101', 'coljac': 'lunar_ascent_rk4_optimizer_eval_N16384_nt101',
        'picard': 'lotka_volterra_rk4_picard_adjoint_N1M_nt101'}[which]
wl = dict(bench.WORKLOADS[name])
if which in ('pde', 'pde32', 'pdetr'):
    wl.update(nt=17, nc=4)            # 16 steps, 4 segments: same kernels, short launch list
if which == 'picard':
    wl.update(N=1 << 17)              # 131,072 trajectories: 16k CTAs, same kernel


class A:
    newton_tol, e2e_slices, world, no_e2e_graph, e2e_fractions, e2e_depth = 1e-7, 2, 1, True, '', 1


run = bench.Runner(wl, name, 0, torch.device('cuda', 0), A)
for _ in range(3):
    run.forward()
    run.adjoint()
torch.cuda.synchronize()
print('done', which)
