#!/usr/bin/env python
"""Headline benchmark: state-steps/s (forward + adjoint) of BASELINE.json config C2
(Van der Pol, Gauss-Legendre 6 TimeMarching + adjoint, 65,536 parameter sets per GPU, nt = 40).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

One JSON line on rank 0.  A "step" = one forward solve + one reverse sweep over the whole batch
(C2: gradient of sum_b J_b(tf) w.r.t. every control value u[t, b] and every step size h[k]).
  value : state-steps/s, inputs resident in HBM, CUDA-event time of the kernels, L2 flushed
          between steps, max over ranks.  state-steps = N * d * n.
  e2e   : same metric through the plan API with HOST (pinned) inputs: H2D of the inputs, solve,
          reverse sweep, D2H of the objective and the gradients inside the timed region.
  roofline / cpu_baseline : see DESIGN.md "Measurement".
Under torchrun every rank owns its own batch (weak scaling); the only collective is the
all-reduce of the gradients of batch-shared inputs (a few hundred scalars).
Other workloads (--workload) cover the remaining BASELINE configs; the default is the headline.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# which CUDA library the run loads: the in-tree build unless OZONE_B200_LIB points elsewhere (A/B experiments only)
LIB_NOTE = os.environ.get('OZONE_B200_LIB') or 'ozone_b200/liboz_b200.so (in-tree build)'

# name: dict(kind, rhs, method, N per GPU, nt, horizon, dtype)
WORKLOADS = {
    'van_der_pol_gl6_tm_adjoint_N65536_nt40': dict(kind='tm', rhs='van_der_pol', method='GaussLegendre6', N=65536, nt=40, T=15.0),
    'van_der_pol_gl4_tm_adjoint_N65536_nt40': dict(kind='tm', rhs='van_der_pol', method='GaussLegendre4', N=65536, nt=40, T=15.0),
    'van_der_pol_gl2_tm_adjoint_N65536_nt40': dict(kind='tm', rhs='van_der_pol', method='GaussLegendre2', N=65536, nt=40, T=15.0),
    'lotka_volterra_rk4_tm_adjoint_N1M_nt101': dict(kind='tm', rhs='lotka_volterra', method='RK4', N=1 << 20, nt=101, T=0.9),
    'lotka_volterra_rk4_tm_adjoint_N1M_nt101_fp32': dict(kind='tm', rhs='lotka_volterra', method='RK4', N=1 << 20, nt=101, T=0.9, fp32=True),
    'lotka_volterra_rk4_picard_adjoint_N1M_nt101': dict(kind='picard', rhs='lotka_volterra', method='RK4', N=1 << 20, nt=101, T=0.9),
    'lotka_volterra_rk4_picard_adjoint_N1M_nt101_fp32': dict(kind='picard', rhs='lotka_volterra', method='RK4', N=1 << 20, nt=101, T=0.9, fp32=True),
    'lotka_volterra_rk4_picard_jacobi_adjoint_N1M_nt101': dict(kind='picard', rhs='lotka_volterra', method='RK4', N=1 << 20, nt=101, T=0.9, jacobi=True),
    'diffusion_reaction_rk4_checkpoints_nx1M_n1024': dict(kind='grid', rhs='diffusion_reaction', method='RK4', N=1, nt=1025, nx=1000000, nc=32),
    'diffusion_reaction_rk4_checkpoints_nx1M_n1024_fp32': dict(kind='grid', rhs='diffusion_reaction', method='RK4', N=1, nt=1025, nx=1000000, nc=32, fp32=True),
    # the same solve with the whole history kept (approach TimeMarching, 8.2 GB): what the checkpoint recompute costs
    # the same PDE given as a Python callback written like the reference's (pde_control.py:11-81) and compiled by trace_grid.py
    'diffusion_reaction_rk4_checkpoints_nx1M_n1024_traced': dict(kind='grid', rhs='diffusion_reaction', method='RK4', N=1, nt=1025, nx=1000000, nc=32, traced=True),
    'diffusion_reaction_rk4_tm_history_nx1M_n1024': dict(kind='grid', rhs='diffusion_reaction', method='RK4', N=1, nt=1025, nx=1000000, nc=0),
    # the same grid cut into spatial slabs over the ranks (strong scaling: the 10^6 cells are shared out)
    'diffusion_reaction_rk4_checkpoints_nx1M_n1024_sharded': dict(kind='grid', rhs='diffusion_reaction', method='RK4', N=1, nt=1025, nx=1000000, nc=32, sharded=True),
    'lunar_ascent_rk4_collocation_N16384_nt101': dict(kind='colloc', rhs='lunar_ascent', method='RK4', N=16384, nt=101, T=0.1),
    # one optimiser evaluation of config C5: constraints c (residual kernel) + the factored constraint Jacobian dc
    # (f, df/dy, df/dp at every stage node: optimize.CollocationBlocks)
    'lunar_ascent_rk4_optimizer_eval_N16384_nt101': dict(kind='coljac', rhs='lunar_ascent', method='RK4', N=16384, nt=101, T=0.1),
    'lunar_ascent_gl6_collocation_N16384_nt101': dict(kind='colloc', rhs='lunar_ascent', method='GaussLegendre6', N=16384, nt=101, T=0.1),
}
DEFAULT_WORKLOAD = 'van_der_pol_gl6_tm_adjoint_N65536_nt40'
STATE_SIZE = {'van_der_pol': 3, 'lotka_volterra': 2, 'lunar_ascent': 7}     # packed state components per trajectory
# compact entries of the default bench line (one short run each): the other BASELINE configs, driver-timed
OTHER_WORKLOADS = ('lotka_volterra_rk4_tm_adjoint_N1M_nt101', 'diffusion_reaction_rk4_checkpoints_nx1M_n1024',
                   'diffusion_reaction_rk4_checkpoints_nx1M_n1024_traced', 'lunar_ascent_rk4_collocation_N16384_nt101')


def workload_config(name, wl, world, newton_tol):
    """The `config` object BOTH arms print, identical for the same command line, so the two lines compare field by field.
    What is specific to an arm (sampling of the CPU run, solver statistics of the GPU run) is printed beside it."""
    d = wl.get('nx') if wl['kind'] == 'grid' else STATE_SIZE[wl['rhs']]
    if wl['kind'] == 'grid':
        par = (f'spatial slabs x{world} (halo exchange between ranks)' if wl.get('sharded') else
               'replicas only (one grid per GPU); the spatial shard is timed by the ..._sharded workload')
    else:
        par = f'trajectory-batch x{world}'
    return {'workload': name, 'rhs': wl['rhs'], 'method': wl['method'], 'approach': wl['kind'],
            'trajectories_per_gpu': wl['N'], 'num_times': wl['nt'], 'state_size': d, 'newton_tol': newton_tol,
            'l2': 'GPU arm: flushed (256 MiB write) between timed steps; e2e with several solves in flight: every step\'s inputs arrive over PCIe', 'parallelism': par}


ASCENT_IC = np.array([0.061663494666509, 0.055249439788994, -1.004110522066614, -0.465976472038572,
                      -0.416909510597800, -0.084973690970470, 11846.599614336959 / 15103.])


def pde_control_callback(ozone_vars, dx, nx, beta_t, beta_u, gamma):
    """The reference's PDE right-hand side as a user would write it against ozone_b200.ops (pde_control.py:11-81)."""
    from ozone_b200 import ops
    heaviside = np.zeros((nx, 1))
    heaviside[round(0.3 * nx):round(0.7 * nx)] = 1.0
    x = ops.reshape(ozone_vars.states['x'][0, :], (nx, 1))
    u = ozone_vars.dynamic_parameters['u']
    zeros = np.zeros((1, 1))
    lap = (ops.concatenate((zeros, x[:-1, :])) + (-2.0) * x + ops.concatenate((x[1:, :], zeros))) / dx ** 2
    reaction = beta_t * (np.exp(1) ** (-gamma * ((1 + x) ** (-1)))) + (-1 * beta_t * np.exp(-gamma))
    transfer = ops.expand(ops.reshape(u[0], (1,)), (nx, 1)) * (beta_u * heaviside) + (-beta_u) * x
    ozone_vars.d_states['x'] = (lap + reaction + transfer).reshape(1, -1)
    xs, us = ops.reshape(x, (nx,)), ops.reshape(u[0, :], (1,))
    ozone_vars.field_outputs['J_space'] = ((100.0 / nx) * ops.sum(xs * xs) + 20.0 * (us * us)).reshape(1, 1)


def synth_inputs(wl, N, seed):
    """Seeded synthetic batch of the named shape."""
    rng = np.random.default_rng(seed)
    rhs, nt = wl['rhs'], wl['nt']
    if rhs == 'diffusion_reaction':
        nx = wl['nx']
        dx = np.pi / nx
        return dict(y0=0.5 * np.ones((nx, 1)), h=np.full(nt - 1, 0.4 * dx * dx),
                    p=rng.uniform(-0.6, 0.0, size=(nt, 1, 1)), th=None)
    h = np.full(nt - 1, wl['T'] / (nt - 1))
    if rhs == 'van_der_pol':
        return dict(y0=np.tile(np.array([[1.0], [1.0], [0.0]]), (1, N)), h=h,
                    p=rng.uniform(-0.75, 1.0, size=(nt, 1, N)), th=None)
    if rhs == 'lotka_volterra':
        t = np.arange(nt)
        return dict(y0=rng.uniform(1.5, 2.5, size=(2, N)), h=h, p=(1.0 + t / nt / 5.0).reshape(nt, 1, 1),
                    th=rng.uniform(0.4, 0.6, size=(1, N)))
    if rhs == 'lunar_ascent':
        return dict(y0=np.tile(ASCENT_IC[:, None], (1, N)), h=h, p=rng.uniform(-1.0, 1.0, size=(nt, 3, N)), th=None)
    raise KeyError(rhs)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--id={self.index}', f'--query-gpu={self.Q}',
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), r[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'samples': len(sm), 'reasons': sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (numpy port of the reference algorithm) on the host cores
def _oracle_chunk(args):
    os.environ['OPENBLAS_NUM_THREADS'] = '1'      # as the reference's scripts do (ozone/tests/run_ode.py:2-4)
    wl, case = args
    from oracle import rhs as R, integrators as I, adjoint as Adj
    kind, method = wl['kind'], wl['method']
    if wl['rhs'] == 'diffusion_reaction':
        rhs = R.DiffusionReaction(case['y0'].shape[0])
        out = I.time_marching(rhs, method, case['y0'], case['h'], case['p'], case['th'])
        g = Adj.tm_adjoint(rhs, method, case['y0'], case['h'], case['p'], case['th'], history=out['history'],
                           field_bar=np.ones_like(out['field']))
        return float(out['field'].sum()), float(np.abs(g['p']).sum())
    rhs = R.REGISTRY[wl['rhs']]()
    if kind in ('colloc', 'coljac'):
        from oracle.tableaux import tableau
        n, s = len(case['h']), tableau(method)[0].shape[0]
        Y, y = I.collocation_guess(case['y0'], n, s)
        c1, c2 = I.collocation_residual(rhs, method, case['y0'], case['h'], Y, y, case['p'], case['th'])
        g = Adj.colloc_vjp(rhs, method, case['y0'], case['h'], Y, y, case['p'], case['th'], np.ones_like(c1), np.ones_like(c2))
        return float(c1.sum()), float(np.abs(g['p']).sum())
    if kind == 'picard':
        out = I.picard(rhs, method, case['y0'], case['h'], case['p'], case['th'], tol=1e-12,
                       state_blocks=[slice(0, 1), slice(1, 2)])
    else:
        out = I.time_marching(rhs, method, case['y0'], case['h'], case['p'], case['th'])
    fb = np.zeros_like(out['final'])
    fb[-1] = 1.0
    g = Adj.tm_adjoint(rhs, method, case['y0'], case['h'], case['p'], case['th'], history=out['history'], final_bar=fb)
    return float(out['final'][-1].sum()), float(np.abs(g['p']).sum())


def cpu_port_throughput(wl, cores, units_per_core, repeats=1):
    """Time the oracle on `cores` processes, each working on `units_per_core` trajectories (or, for
    the grid workload, a grid of that many cells and 64 steps).  Returns (state-steps/s, sample, seconds)."""
    import multiprocessing as mp
    jobs = []
    w = dict(wl)
    if wl['rhs'] == 'diffusion_reaction':
        w.update(nx=units_per_core, nt=65)
        d, n = units_per_core, 64
        per_job = d * n
        what = f'{cores} processes x one {units_per_core}-cell grid, 64 steps'
    else:
        d = STATE_SIZE[wl['rhs']]
        n = wl['nt'] - 1
        per_job = units_per_core * d * n
        what = f'{cores} processes x {units_per_core} trajectories'
    for c in range(cores):
        jobs.append((w, synth_inputs(w, units_per_core, seed=100 + c)))
    ctx = mp.get_context('fork')
    best = float('inf')
    with ctx.Pool(cores) as pool:
        small = dict(w)
        if wl['rhs'] == 'diffusion_reaction':
            small.update(nx=64)
        pool.map(_oracle_chunk, [(small, synth_inputs(small, 8, seed=1))] * cores)     # warm-up (imports, page-in)
        for _ in range(repeats):
            t0 = time.perf_counter()
            pool.map(_oracle_chunk, jobs)
            best = min(best, time.perf_counter() - t0)
    sample = (f'{what}, numpy oracle port of the reference algorithm (OPENBLAS_NUM_THREADS=1 each), '
              f'{wl["kind"]} forward + adjoint, best of {repeats}')
    return cores * per_job / best, sample, best


CPU_SAMPLE = {'tm': 4096, 'picard': 256, 'colloc': 2048, 'coljac': 2048, 'grid': 4096}       # per core, ~10-30 s of CPU work in total
CPU_SAMPLE_REF = {'tm': 1024, 'picard': 64, 'colloc': 512, 'coljac': 512, 'grid': 2048}      # per step of --impl reference


def run_reference(args, rank, world):
    """--impl reference: the reference algorithm on the host cores.  The reference itself cannot run
    (csdl_alpha is an un-vendored, uninstallable dependency), so this times the oracle port."""
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    per_core = args.cpu_traj_per_core or CPU_SAMPLE_REF[wl['kind']]
    vals = []
    for _ in range(max(args.warmup, 0)):
        cpu_port_throughput(wl, cores, max(per_core // 4, 8))
    t_all = 0.0
    for _ in range(args.steps):
        v, sample, dt = cpu_port_throughput(wl, cores, per_core)
        vals.append(v)
        t_all += dt
    value = float(np.mean(vals))
    line = {
        'impl': 'reference', 'metric': 'state-steps/sec (fwd+adjoint)', 'value': value, 'unit': 'state-steps/s',
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * t_all / args.steps,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': workload_config(args.workload, wl, world, args.newton_tol),
        'sampling': f'same workload, sampled: each step is a bounded sample ({sample})',
        'cpu_baseline': {'value': value, 'unit': 'state-steps/s', 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': 'state-steps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
class ShardedGridRunner:
    """Config C3 with the grid cut into one slab per rank (ozone_b200/parallel.py): strong scaling.  The forward sweep
    and the reverse sweep are each captured as ONE CUDA graph (kernels + the peer-memory halo kernels, nothing else);
    the grid-wide sums (J, h_bar, u_bar) are all-reduced once per step, inside the timed region."""

    def __init__(self, wl, name, rank, world, dev, args):
        import torch
        import torch.distributed as dist
        from ozone_b200 import engine, parallel
        from ozone_b200.glms import get_integration_method
        self.torch, self.engine, self.dist = torch, engine, dist
        self.wl, self.kind, self.world = wl, 'grid', world
        self.dtype = torch.float32 if wl.get('fp32') else torch.float64
        self.w = 4 if wl.get('fp32') else 8
        nx, nt = wl['nx'], wl['nt']
        self.n = nt - 1
        self.d = nx
        self.units = nx * self.n                      # the WHOLE grid: total work is fixed as ranks are added
        case = synth_inputs(wl, 1, seed=0)            # the same problem on every rank
        m = get_integration_method(wl['method'])
        ni = self.n // wl['nc']
        cfg = (np.pi / nx, 16.0, 2.0, 2.0, float(round(0.3 * nx)), float(round(0.7 * nx)))
        t = lambda a: torch.tensor(np.ascontiguousarray(a), dtype=self.dtype, device=dev)
        self.channel = parallel.make_channel(m, nx, ni, self.dtype, dev) if world > 1 else None
        self.shard = parallel.GridShard(m, nx, self.n, ni, rank, world, t(case['y0'][:, 0]), t(case['h']), t(case['p'][:, 0, 0]),
                                        cfg, self.dtype, dev, channel=self.channel)
        self.spec = self.shard._spec(0, False)
        self.pipe, self.plan = None, None
        self.host, self.e2e_inputs, self.h2d, self.d2h = {}, (), 0, 0
        (K, KR), ni_ = engine.grid_fusion(self.shard.slab.n_ext), ni
        groups = lambda steps, k: -(-steps // k)
        nseg = self.shard.nseg
        self.launches = nseg * (groups(ni_, K) + 2 + groups(ni_ - 1, K) + groups(ni_, KR) + 2
                                + (4 if world > 1 else 0))
        self.graphs = None
        self.packed = torch.zeros(1 + self.n + self.n + 1, dtype=self.dtype, device=dev)
        if not args.no_shard_graph:
            self._capture()

    def _fwd(self):
        sh = self.shard
        for seg in range(sh.nseg):
            sh.forward_segment(seg)
            sh.exchange(sh.ckpts[seg + 1])

    def _adj(self):
        sh = self.shard
        sh.start_adjoint(1.0)
        for seg in reversed(range(sh.nseg)):
            sh.adjoint_segment(seg)
            sh.exchange(sh.lam_cur)
        if sh.channel is not None:
            sh.channel.advance(self.torch.cuda.current_stream().cuda_stream)
        r = sh.results()
        self.packed[0:1].copy_(r['field'].reshape(-1))
        self.packed[1:1 + self.n].copy_(r['h_bar'])
        self.packed[1 + self.n:].copy_(r['u_bar'].reshape(-1))

    def _capture(self):
        torch = self.torch
        self._fwd(); self._adj()                      # warm-up outside the capture
        torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        gf, ga = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(gf):
            self._fwd()
        with torch.cuda.graph(ga):
            self._adj()
        self.graphs = (gf, ga)

    def forward(self):
        self.graphs[0].replay() if self.graphs else self._fwd()

    def adjoint(self):
        self.graphs[1].replay() if self.graphs else self._adj()
        if self.world > 1:
            self.dist.all_reduce(self.packed)         # J, h_bar, u_bar: sums over the slabs

    def shared_grad(self):
        return None                                   # reduced inside adjoint()

    def e2e(self):
        self.forward(); self.adjoint()

    def e2e_readback(self):
        pass

    def check(self):
        if self.channel is not None:
            self.channel.check()

    def algorithmic_bytes(self):
        w, n = self.w, self.n
        d = self.shard.slab.own_hi - self.shard.slab.own_lo
        return 2 * w * d * n, 5 * w * d * n


class Runner:
    """One workload on one rank: device-resident step (fwd, adj timed separately) and e2e step."""

    def __init__(self, wl, name, rank, dev, args, e2e=True):
        import torch
        from ozone_b200 import engine
        from ozone_b200.glms import get_integration_method
        self.torch, self.engine = torch, engine
        self.wl, self.kind = wl, wl['kind']
        self.dtype = torch.float32 if wl.get('fp32') else torch.float64
        self.w = 4 if wl.get('fp32') else 8
        N, nt = wl['N'], wl['nt']
        self.n = nt - 1
        case = synth_inputs(wl, N, seed=rank)
        self.p_shared = case['p'].shape[-1] == 1
        extra = {}
        if self.kind == 'grid':
            nx = wl['nx']
            extra = dict(d_override=nx, rhs_cfg=(np.pi / nx, 16.0, 2.0, 2.0, float(round(0.3 * nx)), float(round(0.7 * nx))),
                         num_inner=(self.n // wl['nc']) if wl.get('nc') else 0, store_history=not wl.get('nc'))
        if self.kind == 'picard':
            extra = dict(state_blocks=(0, 2) if wl.get('jacobi') else (0, 1, 2))
        traced = None
        if wl.get('traced'):      # the callback goes through the symbolic grid tracer; its coefficient field travels as th
            from ozone_b200.trace_grid import TracedGridRHS
            nx = wl['nx']
            traced = TracedGridRHS(pde_control_callback, (), dict(dx=np.pi / nx, nx=nx, beta_t=16.0, beta_u=2.0, gamma=2.0), 'x', nx, ['u'])
            extra.pop('rhs_cfg')
            case['th'] = traced.tracer.coefficient_table().reshape(-1)
        self.spec = spec = engine.SolveSpec(
            rhs_name='traced' if traced is not None else wl['rhs'], traced=traced, method=get_integration_method(wl['method']),
            N=N, n=self.n, dtype=self.dtype,
            h_shared=True, p_shared=self.p_shared, th_shared=traced is not None, newton_tol=args.newton_tol,
            **({'store_history': True} if self.kind != 'grid' else {}), **extra)
        self.d = spec.d
        self.units = N * self.d * self.n
        want = ('h', 'p', 'th') if wl['rhs'] == 'lotka_volterra' else (('y0', 'h', 'p') if self.kind == 'grid' else ('h', 'p'))
        cot = ('field',) if self.kind == 'grid' else ('final',)
        self.plan = plan = engine.TmPlan(spec, device=dev, grads=want, cotangents=cot)
        t = lambda a: torch.tensor(np.ascontiguousarray(a), dtype=self.dtype)
        self.host = {'y0': t(case['y0']).pin_memory(), 'h': t(case['h']).pin_memory(),
                     'p': t(case['p'][:, :, 0] if self.p_shared else case['p']).pin_memory()}
        if case['th'] is not None:
            self.host['th'] = t(case['th']).pin_memory()
        for k, v in self.host.items():
            getattr(plan, k).copy_(v)
        if self.kind == 'grid':
            plan.field_bar.fill_(1.0)
        else:
            plan.final_bar.zero_()
            plan.final_bar[-1].fill_(1.0) if wl['rhs'] == 'van_der_pol' else plan.final_bar.fill_(1.0)
        self.host_out = {k: torch.empty_like(g, device='cpu').pin_memory() for k, g in plan.grad.items() if g is not None}
        self.obj = plan.field if self.kind == 'grid' else plan.final[-1]
        self.host_obj = torch.empty_like(self.obj, device='cpu').pin_memory()
        self.e2e_inputs = tuple(k for k in ('y0', 'h', 'p', 'th') if k in self.host)      # every input crosses the link each step
        self.h2d = sum(self.host[k].numel() * self.w for k in self.e2e_inputs)
        self.d2h = sum(v.numel() * self.w for v in self.host_out.values()) + self.host_obj.numel() * self.w
        self.launches = plan.launches_forward + plan.launches_adjoint
        self.pipe = None
        if self.kind == 'tm' and e2e:      # e2e leg: host-resident batch in sub-batches on their own streams (copy / compute overlap)
            fr = [float(x) for x in args.e2e_fractions.split(',')] if args.e2e_fractions else None
            self.pipe = engine.PipelinedTmSolve(spec, slices=args.e2e_slices, device=dev, grads=want, cotangents=cot,
                                                upload=self.e2e_inputs, fractions=fr)
            # steps of the end-to-end leg are independent solves of host-resident batches: `depth` of them are in flight, each with
            # its own device + pinned buffers and streams, so that step k + 1 uploads while step k computes and downloads
            depth = max(1, int(getattr(args, 'e2e_depth', 1)))
            self.queue = engine.SolveStream(spec, depth=depth, device=dev, slices=args.e2e_slices, grads=want, cotangents=cot,
                                            upload=self.e2e_inputs, fractions=fr)
            self.pipes = [self.pipe] + self.queue.pipes[1:]      # slot 0 is the solve that is also timed alone
            self.queue.pipes = self.pipes
            for pipe in self.pipes:
                pipe.set_host_inputs(**self.host)
                if args.world > 1:
                    import torch.distributed as dist
                    pipe.reduce_shared = dist.all_reduce
                for pl in pipe.plans:
                    pl.final_bar.copy_(plan.final_bar[:, :pl.spec.N])
            self.h2d, self.d2h = self.pipe.bytes_per_run()
            self.e2e_launches = self.pipe.launches
            self.e2e_graph = False
            if not args.no_e2e_graph:
                try:                       # replay the slice pipeline as one CUDA graph (host enqueue cost off the path)
                    for pipe in self.pipes:
                        pipe.capture()
                    self.e2e_graph = True
                except Exception as exc:   # noqa: BLE001 -- fall back to eager enqueue, and say so in the bench line
                    for pipe in self.pipes:
                        pipe.graph = None
                    self.e2e_graph_error = repr(exc)[:200]
                    torch.cuda.synchronize()
        if self.kind == 'picard':
            self.launches = 1 + plan.launches_adjoint
        if self.kind in ('colloc', 'coljac'):
            s = spec.method.num_stages
            rng = np.random.default_rng(1000 + rank)
            dev_t = lambda a: torch.tensor(a, dtype=self.dtype, device=dev)
            self.Y = dev_t(np.broadcast_to(case['y0'], (self.n, s, self.d, N)).copy())      # collections.py:129-133
            self.y = dev_t(np.broadcast_to(case['y0'], (self.n, self.d, N)).copy())
            self.mu1 = dev_t(rng.standard_normal((self.n, s, self.d, N)))
            self.mu2 = dev_t(rng.standard_normal((self.n, self.d, N)))
            self.launches = 2 + 1       # residual, vjp, h_bar reducer
            self.host_colloc = None

    # -- pieces timed separately on the device
    def forward(self):
        if self.kind == 'picard':
            pl = self.plan
            self.pic = self.engine.picard_forward(self.spec, pl.y0, pl.h, pl.p, pl.th, tol=1e-12 if self.w == 8 else 1e-4)
        elif self.kind in ('colloc', 'coljac'):
            pl = self.plan
            self.c = self.engine.colloc_residual(self.spec, pl.y0, pl.h, pl.p, pl.th, self.Y, self.y)
        else:
            self.plan.forward()

    def adjoint(self):
        if self.kind == 'picard':
            pl = self.plan
            self.g = self.engine.tm_adjoint(self.spec, pl.y0, pl.h, pl.p, pl.th, hist=self.pic['hist'], final_bar=pl.final_bar,
                                            need=('h', 'p', 'th'))
        elif self.kind == 'coljac':
            pl, sp = self.plan, self.spec
            n, st, d, N, P = sp.n, sp.method.num_stages, sp.d, sp.N, sp.info['P']
            self.jac = self.engine.rhs_jac_eval(sp, self.Y.reshape(n * st, d, N), pl.p, pl.th, compact=True, kind=self.engine.KIND_COLLOC,
                                                stage_params=True)
            self.g = {'h': None}
        elif self.kind == 'colloc':
            pl = self.plan
            self.g = self.engine.colloc_vjp(self.spec, pl.y0, pl.h, pl.p, pl.th, self.Y, self.y, self.mu1, self.mu2,
                                            need=('h', 'p'))
        else:
            self.plan.adjoint()

    def shared_grad(self):
        if self.kind in ('picard', 'colloc', 'coljac'):
            return self.g['h']
        return self.plan.grad['h']

    def e2e(self):
        if self.pipe is not None:
            self.pipe.run()
            return
        pl = self.plan
        for k in self.e2e_inputs:
            getattr(pl, k).copy_(self.host[k], non_blocking=True)
        self.forward()
        self.adjoint()

    def e2e_async(self, k):
        """Step k of a run of independent end-to-end solves (engine.SolveStream: slot k mod depth)."""
        self.queue.submit()

    def e2e_fork(self):
        self.queue.begin()

    def e2e_join(self):
        self.queue.end()

    def e2e_readback(self):
        if self.pipe is not None:
            return                  # downloads are part of every slice's stream
        if self.kind in ('picard', 'colloc', 'coljac'):
            src = self.pic['final'][-1] if self.kind == 'picard' else self.c[1][-1, 0]
            self.host_obj.copy_(src, non_blocking=True)
            for k, v in self.host_out.items():
                if self.g.get(k) is not None:
                    v.copy_(self.g[k], non_blocking=True)
        else:
            self.host_obj.copy_(self.obj, non_blocking=True)
            for k, v in self.host_out.items():
                v.copy_(self.plan.grad[k], non_blocking=True)

    def algorithmic_bytes(self):
        """Compulsory HBM bytes of one forward and one adjoint call (DESIGN.md 'Measurement')."""
        w, N, n, d = self.w, self.wl['N'], self.n, self.d
        P = self.spec.info['P']
        pN = 1 if self.p_shared else N
        if self.kind == 'grid':       # per cell-step: fwd 2w; replay 2w; adjoint step 3w  
            return 2 * w * d * n, (5 if self.wl.get('nc') else 3) * w * d * n
        if self.kind == 'coljac':
            s = self.spec.method.num_stages
            fwd = w * (2 * n * s * d * N + 2 * n * d * N + (n + 1) * P * pN)          # read Y, y, p; write c1, c2
            jp, jpp = self.engine.rhs_jac_pattern(self.spec, self.engine.KIND_COLLOC)
            adj = w * N * (n * s * (d + d + int(jp.sum()) + int(jpp.sum())) + (n + 1) * P)   # read Y and the raw p; write f and the structurally nonzero entries of df/dy, df/dp
            return fwd, adj
        if self.kind == 'colloc':
            s = self.spec.method.num_stages
            fwd = w * (2 * n * s * d * N + 2 * n * d * N + (n + 1) * P * pN)          # read Y, y, p; write c1, c2
            adj = w * (3 * n * s * d * N + 3 * n * d * N + 2 * (n + 1) * P * pN)      # read mu1, mu2, Y, p; write Ybar, ybar, pbar
            return fwd, adj
        info = self.spec.info
        S, tN = info['S'], (1 if self.spec.th_shared else N)
        outs = ((n + 1) * info['nprof'] + info['nfield']) * N                         # profile rows + field sums the RHS declares
        fwd = w * (d * N + (n + 1) * P * pN + S * tN + (n + 1) * d * N + d * N + outs)   # y0, p, th | hist (incl. IC row), final, outputs
        adj = w * (n * d * N + d * N + (n + 1) * P * pN + S * tN + (n + 1) * P * pN + S * tN)   # y_0..y_{n-1}, final_bar, p, th | p_bar, th_bar
        return fwd, adj


FP64_BOUND_NOTE = {
    'tm_implicit': ('implicit Gauss-Legendre + Newton is FP64-ALU bound, not HBM bound (about 1 kflop per 100 compulsory bytes; '
                    'ncu: fp64 pipe ~53 % busy vs DRAM ~12 %): frac is executed FP64 flops over the measured FMA peak, the HBM '
                    'fractions the metric asks for are under roofline.hbm'),
    'grid': ('one exponential per cell and stage: the fused steps are FP64-bound (ncu: fp64 pipe 50-60 %, state L2-resident); HBM '
             'fractions under roofline.hbm'),
    'other': ('this kernel is closer to the FP64 roof than to the HBM one (ncu, profiles/r2c_ncu_summary.txt: explicit reverse sweep '
              'fp64 pipe 78 % busy vs DRAM 26 %; Picard sweeps never leave the chip); HBM fractions under roofline.hbm'),
}
FP64_PEAK_FALLBACK_TFLOPS = 36.3      # oz_fma_probe on this pool's B200 (used only when the live probe is skipped)


def roofline_of(run, workload, fwd_s, adj_s, torch, dev, probe=True):
    """roofline object of the bench line: the dominant kernel against the resource that bounds it.
    HBM-bound families: algorithmic bytes / launch time over the measured copy bandwidth.
    FP64-bound families (implicit time marching, the grid in fp64): executed FP64 flops per launch (ncu count of the
    shipped build, profiles/traffic.json) / launch time over the FMA peak measured in this run; the HBM fractions stay
    available under 'hbm' because the metric asks for them."""
    fwd_b, adj_b = run.algorithmic_bytes()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    peak = float(peaks.get('hbm_gbs', 6650.0))
    names = {'tm': ('tm_adjoint_kernel', 'tm_forward_kernel'), 'picard': ('tm_adjoint_kernel', 'picard_kernel'),
             'grid': ('pde_adjoint_step_kernel (+ replayed pde_step_kernel)', 'pde_step_kernel'),
             'colloc': ('colloc_vjp_kernel', 'colloc_residual_kernel'),
             'coljac': ('rhs_jac_kernel', 'colloc_residual_kernel')}[run.kind]
    traffic = {}
    try:
        traffic = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json'))).get(
            workload.replace('_sharded', '').replace('rk4_tm_history', 'rk4_checkpoints'), {})
    except Exception:
        pass

    # profiles/traffic.json holds ncu counts PER CAPTURED LAUNCH; one timed forward / reverse call of the bench is many
    # launches for the grid (fused steps) and a larger batch for Picard (captured at 131,072 trajectories)
    def per_call(table, name):
        key = name.split(' ')[0]
        v = table.get(key)
        if v is None:
            return None
        if run.kind == 'grid':
            nx = run.shard.slab.n_ext if hasattr(run, 'shard') else run.d
            K, KR = run.engine.grid_fusion(nx)
            n = run.n
            ni = n // run.wl['nc'] if run.wl.get('nc') else n               # no checkpoints: one segment, nothing replayed
            nseg = n // ni
            up = lambda a, b: -(-a // b)
            if key == 'pde_step_kernel':
                return v * nseg * up(ni, K)
            step = table.get('pde_step_kernel') or 0
            replay = step * nseg * up(ni - 1, K) if run.wl.get('nc') else 0
            return v * nseg * up(ni, KR) + replay                            # reverse launches + the replayed forward ones
        if run.kind == 'picard' and key == 'picard_kernel':
            return v * run.wl['N'] / 131072.0
        return v

    def hbm_entry(name, nbytes, secs):
        ach = nbytes / secs / 1e9
        return {'kernel': name, 'achieved': ach, 'peak': peak, 'unit': 'GB/s', 'frac': ach / peak,
                'frac_of_nominal_8000': ach / 8000.0, 'traffic': per_call(traffic, name),
                'algorithmic_bytes_per_launch': nbytes, 'avg_launch_ms': secs * 1e3}

    adj_e, fwd_e = hbm_entry(names[0], adj_b, adj_s), hbm_entry(names[1], fwd_b, fwd_s)
    dominant, other = (adj_e, fwd_e) if adj_s >= fwd_s else (fwd_e, adj_e)
    fma = {}
    if probe:
        try:
            fma = {'f64': run.engine.measure_fma_peak(torch.float64, dev), 'f32': run.engine.measure_fma_peak(torch.float32, dev)}
        except Exception as exc:                                            # never let the probe break the bench line
            fma = {'error': str(exc)}
    implicit = run.kind == 'tm' and not run.spec.method.explicit
    flops = traffic.get('fp64_flops_per_launch', {})
    f64_peak = fma.get('f64') or FP64_PEAK_FALLBACK_TFLOPS
    roof = None
    # the dominant kernel is reported against the roof it is closer to: executed FP64 flops over the FMA peak, or
    # algorithmic bytes over the copy bandwidth
    fl = per_call(flops, dominant['kernel']) if run.w == 8 else None
    fp64_bound = bool(fl) and (fl / (dominant['avg_launch_ms'] * 1e-3) / 1e12 / f64_peak) > dominant['frac']
    if fp64_bound:
        if fl:
            t = fl / (dominant['avg_launch_ms'] * 1e-3) / 1e12
            roof = {'bound': 'fp64', 'kernel': dominant['kernel'], 'achieved': t, 'peak': f64_peak, 'unit': 'TFLOP/s',
                    'frac': t / f64_peak, 'traffic': dominant['traffic'], 'flops_per_launch': fl,
                    'avg_launch_ms': dominant['avg_launch_ms'], 'flops_source': flops.get('source'),
                    'peak_source': 'oz_fma_probe (register-resident DFMA loop) timed in this run' if fma.get('f64')
                    else 'oz_fma_probe at capture time (profiles/traffic.json)',
                    'note': FP64_BOUND_NOTE['tm_implicit' if implicit else ('grid' if run.kind == 'grid' else 'other')]}
            fo = per_call(flops, other['kernel'])
            if fo:
                to = fo / (other['avg_launch_ms'] * 1e-3) / 1e12
                roof['other_kernel'] = {'kernel': other['kernel'], 'achieved': to, 'peak': f64_peak, 'unit': 'TFLOP/s',
                                        'frac': to / f64_peak, 'avg_launch_ms': other['avg_launch_ms']}
    if roof is None:
        roof = dict({'bound': 'hbm'}, **dominant)
        roof.update({'peak_source': 'MEASURED_PEAKS.json hbm_gbs (burst copy)' if peaks else 'fallback 6650 GB/s',
                     'other_kernel': other,
                     'note': 'HBM-bound by this measure (algorithmic bytes over the measured copy bandwidth exceed the FP64 fraction)'})
    roof['hbm'] = {'forward': fwd_e, 'reverse': adj_e, 'peak_source': 'MEASURED_PEAKS.json hbm_gbs (burst copy)' if peaks
                   else 'fallback 6650 GB/s'}
    roof['traffic_source'] = traffic.get('source')
    if fma:
        roof['fma_peak_measured_tflops'] = fma
    return roof


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument('--cpu-traj-per-core', type=int, default=None,
                    help='CPU sample size per core (defaults: CPU_SAMPLE / CPU_SAMPLE_REF tables)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--shared-sum', default='nccl', choices=['peer', 'nccl'], help='how gradients of batch-shared inputs are summed over ranks')
    ap.add_argument('--no-shard-graph', action='store_true', help='sharded grid: enqueue eagerly instead of replaying CUDA graphs')
    ap.add_argument('--no-other-workloads', action='store_true', help='skip the compact other_workloads entries of the default line')
    ap.add_argument('--other-steps', type=int, default=5)
    ap.add_argument('--trajectories', type=int, default=None, help='override the batch size per GPU (experiments only: the bench line names it)')
    ap.add_argument('--newton-tol', type=float, default=1e-7)
    ap.add_argument('--e2e-slices', type=int, default=4, help='sub-batches (streams) of the end-to-end leg')
    ap.add_argument('--e2e-fractions', default='', help='relative sub-batch sizes of the end-to-end leg, e.g. 4,3,2,1 (overrides --e2e-slices)')
    ap.add_argument('--e2e-depth', type=int, default=3,
                    help='independent end-to-end solves in flight (own buffers and streams each); 1 = each step timed alone')
    ap.add_argument('--no-e2e-graph', action='store_true', help='enqueue the end-to-end pipeline eagerly (no CUDA graph)')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else args.warmup

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    args.world = world
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if args.impl == 'reference':
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: ozone_b200 has no CPU path')
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)

    wl = dict(WORKLOADS[args.workload])
    if args.trajectories:
        wl['N'] = args.trajectories
    sharded = bool(wl.get('sharded'))
    run = ShardedGridRunner(wl, args.workload, rank, world, dev, args) if sharded else Runner(wl, args.workload, rank, dev, args)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # Gradients of batch-shared inputs (h: 39 scalars for C2) are summed over the ranks.
    #   peer (default): one small kernel on the compute stream -- every rank stores its row sums into every peer's
    #       symmetric-memory arena and adds up the world x rows values it received, in rank order (ozone_b200/peer.py,
    #       csrc/oz_peer.cu): ~2 NVLink latencies instead of an NCCL launch (~12-26 us at 2-8 ranks in round 1).
    #   nccl: dist.all_reduce on a side stream, overlapping the L2 flush and the next forward solve; the next reverse
    #       sweep waits for it and the LAST one is inside the timed total (tail below).
    comm = torch.cuda.Stream(device=dev) if world > 1 else None
    pending = [None]
    peer_sum = [None]
    shared_mode = ['none (1 GPU)' if world == 1 else args.shared_sum]
    if world > 1 and args.shared_sum == 'peer' and not sharded:
        try:
            from ozone_b200 import peer
            g0 = run.shared_grad()
            if g0 is not None:
                arena = peer.PeerArena(1 << 20, dev)
                peer_sum[0] = peer.PeerSum(arena, g0.numel(), g0.dtype)
                shared_mode[0] = 'peer: oz_peer_sum kernel over symmetric memory, on the compute stream'
        except Exception as exc:                 # noqa: BLE001 -- e.g. no symmetric memory on this box: use NCCL
            peer_sum[0] = None
            shared_mode[0] = f'nccl on a side stream (peer arena unavailable: {repr(exc)[:120]})'
    elif world > 1:
        shared_mode[0] = 'nccl on a side stream, overlapping the next forward solve; the last one is inside the timed total'

    def allreduce_shared(r, e2e=False):
        if world == 1:
            return
        if e2e and r.pipe is not None:        # PipelinedTmSolve all-reduces its summed shared gradients itself
            return
        g = r.shared_grad()
        if g is None:
            return
        if peer_sum[0] is not None and r is run:
            peer_sum[0](g, g, torch.cuda.current_stream().cuda_stream)
            return
        ready = torch.cuda.Event()
        ready.record()
        comm.wait_event(ready)
        with torch.cuda.stream(comm):
            dist.all_reduce(g)
            done = torch.cuda.Event()
            done.record(comm)
        pending[0] = done

    def join_comm():
        if pending[0] is not None:
            torch.cuda.current_stream().wait_event(pending[0])
            pending[0] = None

    ev = lambda: torch.cuda.Event(enable_timing=True)

    def device_step(r, times):
        flush.fill_(rank & 0xff)
        e0, e1, e2 = ev(), ev(), ev()
        e0.record(); r.forward(); e1.record()
        join_comm()                           # the previous step's all-reduce still owns the shared-gradient buffer
        r.adjoint()
        allreduce_shared(r)                   # peer mode: a kernel on this stream, inside the reverse interval
        e2.record()
        times.append((e0, e1, e2))

    def e2e_step(r, times):
        flush.fill_(rank & 0xff)
        e0, e1 = ev(), ev()
        e0.record()
        r.e2e()
        allreduce_shared(r, e2e=True)
        join_comm()
        r.e2e_readback()
        e1.record()
        times.append((e0, e1))

    def timed_device_steps(r, warmup, steps):
        """-> (sum of forward ms, sum of reverse ms incl. the tail until the last all-reduce is done, wall ms)"""
        for _ in range(warmup):
            device_step(r, [])
        join_comm()
        barrier()
        times = []
        t0 = time.perf_counter()
        for _ in range(steps):
            device_step(r, times)
        tail0, tail1 = ev(), ev()
        tail0.record(); join_comm(); tail1.record()
        barrier()
        wall = time.perf_counter() - t0
        fwd = float(np.sum([a.elapsed_time(b) for a, b, _ in times]))
        adj = float(np.sum([b.elapsed_time(c) for _, b, c in times])) + tail0.elapsed_time(tail1)
        return fwd, adj, wall * 1e3

    for _ in range(args.warmup):
        device_step(run, [])
    join_comm()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    fwd_sum, adj_sum, wall_ms = timed_device_steps(run, 0, args.steps)
    step_ms = fwd_sum + adj_sum
    t_wall = wall_ms * 1e-3

    for _ in range(args.warmup):
        e2e_step(run, [])
    barrier()
    etimes = []
    for _ in range(args.steps):
        e2e_step(run, etimes)
    barrier()
    e2e_single_ms = float(np.sum([a.elapsed_time(b) for a, b in etimes]))     # one solve in flight, L2 flushed between steps
    e2e_depth = len(run.pipes) if getattr(run, 'pipe', None) is not None else 1
    if e2e_depth > 1:
        # throughput of a stream of independent host-resident batches: `depth` solves in flight, timed on the device from
        # before the first upload to after the last download of the K steps (inputs arrive over PCIe every step)
        def e2e_run(steps):
            a, b = ev(), ev()
            a.record()
            run.e2e_fork()
            for k in range(steps):
                run.e2e_async(k)
            run.e2e_join()
            b.record()
            return a, b
        e2e_run(args.warmup)
        torch.cuda.synchronize()
        barrier()
        etimes = [e2e_run(args.steps)]
        barrier()
    clocks = sampler.stop() if rank == 0 else None
    if hasattr(run, 'check'):
        run.check()                            # a halo wait that timed out shows up here, not as a hang
    if peer_sum[0] is not None:
        peer_sum[0].check()
    e2e_ms = float(np.sum([a.elapsed_time(b) for a, b in etimes]))

    plan = run.plan
    extra_cfg = {}
    if run.kind == 'tm':
        extra_cfg['not_converged'] = int((plan.status != 0).sum().item())
        if not run.spec.method.explicit:
            extra_cfg['mean_newton_iters_per_step'] = float(plan.iters.double().mean().item()) / run.n
    if run.kind == 'picard':
        extra_cfg['not_converged'] = int((run.pic['status'] != 0).sum().item())
        extra_cfg['mean_sweeps'] = float(run.pic['iters'].double().mean().item())

    tot = torch.tensor([step_ms, e2e_ms, fwd_sum, adj_sum, e2e_single_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    step_ms, e2e_ms, fwd_tot, adj_tot, e2e_single_ms = tot.tolist()

    units_per_step = run.units if sharded else world * run.units
    value = units_per_step * args.steps / (step_ms * 1e-3)
    e2e_value = units_per_step * args.steps / (e2e_ms * 1e-3)

    # the other BASELINE configs, a few steps each, so that the driver-run record covers more than the headline
    others = []
    if args.workload == DEFAULT_WORKLOAD and not args.no_other_workloads and not args.trajectories:
        for name in OTHER_WORKLOADS:
            owl = dict(WORKLOADS[name])
            try:
                orun = Runner(owl, name, rank, dev, args, e2e=False)
                of, oa, _ = timed_device_steps(orun, 3, args.other_steps)
                t = torch.tensor([of, oa], dtype=torch.float64, device=dev)
                if world > 1:
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                of, oa = t.tolist()
                if rank == 0:
                    r = roofline_of(orun, name, of / args.other_steps * 1e-3, oa / args.other_steps * 1e-3, torch, dev, probe=False)
                    others.append({'workload': name, 'value': world * orun.units * args.other_steps / ((of + oa) * 1e-3),
                                   'unit': 'state-steps/s', 'steps': args.other_steps, 'fwd_ms': of / args.other_steps,
                                   'adj_ms': oa / args.other_steps, 'bound': r['bound'], 'frac': r['frac'],
                                   'hbm_frac_fwd_adj': [r['hbm']['forward']['frac'], r['hbm']['reverse']['frac']],
                                   'parallelism': workload_config(name, owl, world, args.newton_tol)['parallelism']})
                del orun
                torch.cuda.empty_cache()
            except Exception as exc:            # noqa: BLE001 -- an extra must never cost the headline line
                if rank == 0:
                    others.append({'workload': name, 'error': repr(exc)[:200]})

    if rank == 0:
        adj_avg_s = adj_tot / args.steps * 1e-3
        fwd_avg_s = fwd_tot / args.steps * 1e-3
        roof = roofline_of(run, args.workload, fwd_avg_s, adj_avg_s, torch, dev)
        cpu = None
        if not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            per_core = args.cpu_traj_per_core or CPU_SAMPLE[run.kind]
            v, sample, _ = cpu_port_throughput(wl, cores, per_core)
            v1, _, _ = cpu_port_throughput(wl, 1, max(per_core // 4, 8))       # the reference's scripts pin one thread
            cpu = {'value': v, 'unit': 'state-steps/s', 'cores': cores, 'kind': 'port', 'sample': sample,
                   'single_core_value': v1}
        line = {
            'metric': 'state-steps/sec (fwd+adjoint)', 'value': value, 'unit': 'state-steps/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': step_ms / args.steps, 'higher_is_better': True,
            'scaling': 'strong' if sharded else 'weak', 'vs_baseline': None, 'dtype': 'f32' if wl.get('fp32') else 'f64', 'data': 'synthetic',
            'config': workload_config(args.workload, wl, world, args.newton_tol),
            'solver_stats': extra_cfg,
            'shared_gradient_allreduce': shared_mode[0],
            'e2e': {'value': e2e_value, 'unit': 'state-steps/s', 'h2d_bytes_per_step': run.h2d,
                    'd2h_bytes_per_step': run.d2h, 'ms_per_step': e2e_ms / args.steps,
                    'solves_in_flight': e2e_depth,
                    'one_solve_in_flight': {'value': units_per_step * args.steps / (e2e_single_ms * 1e-3),
                                            'ms_per_step': e2e_single_ms / args.steps,
                                            'how': 'each step timed alone (upload -> solve -> download), L2 flushed between steps'},
                    'how': ((f'engine.SolveStream: {e2e_depth} independent solves in flight (own device + pinned buffers each), K steps timed on the device '
                             f'from the first upload to the last download; each solve: ' if e2e_depth > 1 else '') + (f'engine.PipelinedTmSolve: {len(run.pipe.plans) if run.pipe is not None else 1} sub-batches (sizes {[b - a for a, b in run.pipe.bounds] if run.pipe is not None else []}) on their own streams, pinned host buffers, '
                            + ('slice pipeline replayed as one CUDA graph' if getattr(run, 'e2e_graph', False) else
                               'eager enqueue' + (f' (graph capture failed: {run.e2e_graph_error})' if hasattr(run, 'e2e_graph_error') else ''))
                            if run.pipe is not None else 'engine plan API, pinned host buffers, one stream'))},
            'gpu_launches': args.steps * run.launches,
            'roofline': roof, 'cpu_baseline': cpu, 'clocks': clocks,
            'wall_ms_timed_loop': t_wall * 1e3, 'other_workloads': others,
            'library': LIB_NOTE,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
