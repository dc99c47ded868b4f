#!/usr/bin/env python
"""Headline benchmark: state-steps/s (forward + adjoint) of BASELINE.json config C2
(Van der Pol, Gauss-Legendre 6 TimeMarching + adjoint, 65,536 parameter sets per GPU, nt = 40).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload ...]

One JSON line on rank 0.  A "step" = one forward solve + one reverse sweep over the whole batch
(gradient of sum_b J_b(tf) w.r.t. every control value u[t, b] and every step size h[k]).
  value : state-steps/s, inputs resident in HBM, CUDA-event time of the kernels, L2 flushed
          between steps, max over ranks.  state-steps = N * d * n  (d = 3 states, n = 39 steps).
  e2e   : same metric through the plan API with HOST (pinned) inputs: H2D of u and h, solve,
          reverse sweep, D2H of the objective and both gradients inside the timed region.
  roofline / cpu_baseline : see DESIGN.md section "Measurement".
Under torchrun every rank owns its own 65,536 trajectories (weak scaling); the only collective
is the all-reduce of the batch-shared gradient h_bar (39 scalars).
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

WORKLOADS = {
    # name: (rhs, method, N per GPU, nt, horizon)
    'van_der_pol_gl6_tm_adjoint_N65536_nt40': ('van_der_pol', 'GaussLegendre6', 65536, 40, 15.0),
    'van_der_pol_gl4_tm_adjoint_N65536_nt40': ('van_der_pol', 'GaussLegendre4', 65536, 40, 15.0),
    'van_der_pol_gl2_tm_adjoint_N65536_nt40': ('van_der_pol', 'GaussLegendre2', 65536, 40, 15.0),
    'lotka_volterra_rk4_tm_adjoint_N1M_nt101': ('lotka_volterra', 'RK4', 1 << 20, 101, 0.9),
    'lotka_volterra_rk4_tm_adjoint_N1M_nt101_fp32': ('lotka_volterra', 'RK4', 1 << 20, 101, 0.9),
}
DEFAULT_WORKLOAD = 'van_der_pol_gl6_tm_adjoint_N65536_nt40'


def synth_inputs(rhs, N, nt, horizon, seed):
    """Seeded synthetic batch of the named shape."""
    rng = np.random.default_rng(seed)
    h = np.full(nt - 1, horizon / (nt - 1))
    if rhs == 'van_der_pol':
        return dict(y0=np.tile(np.array([[1.0], [1.0], [0.0]]), (1, N)), h=h,
                    p=rng.uniform(-0.75, 1.0, size=(nt, 1, N)), th=None)
    if rhs == 'lotka_volterra':
        t = np.arange(nt)
        return dict(y0=rng.uniform(1.5, 2.5, size=(2, N)), h=h, p=(1.0 + t / nt / 5.0).reshape(nt, 1, 1),
                    th=rng.uniform(0.4, 0.6, size=(1, N)))
    raise KeyError(rhs)


def algorithmic_bytes(rhs_info, N, n, w, p_shared, hist_cot):
    """Compulsory HBM bytes of one forward and one adjoint launch (DESIGN.md 'Measurement')."""
    d, P = rhs_info['d'], rhs_info['P']
    pN = 1 if p_shared else N
    fwd = w * (d * N + (n + 1) * P * pN + (n + 1) * d * N + d * N)      # y0, p | hist (incl. IC row), final
    adj = w * (n * d * N + d * N + (n + 1) * P * pN + (n + 1) * P * pN)  # y_0..y_{n-1}, final_bar, p | p_bar
    if hist_cot:
        adj += w * (n + 1) * d * N
    return fwd, adj


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
def _oracle_chunk(args):
    os.environ['OPENBLAS_NUM_THREADS'] = '1'      # as the reference's scripts do (ozone/tests/run_ode.py:2-4)
    rhs_name, method, case = args
    from oracle import rhs as R, integrators as I, adjoint as Adj
    rhs = R.REGISTRY[rhs_name]()
    out = I.time_marching(rhs, method, case['y0'], case['h'], case['p'], case['th'])
    fb = np.zeros_like(out['final'])
    fb[-1] = 1.0
    g = Adj.tm_adjoint(rhs, method, case['y0'], case['h'], case['p'], case['th'], history=out['history'], final_bar=fb)
    return float(out['final'][-1].sum()), float(np.abs(g['p']).sum())


def cpu_port_throughput(rhs_name, method, nt, horizon, cores, traj_per_core, repeats=1):
    """Time the oracle (numpy port of the reference algorithm) on `cores` processes, each solving
    and differentiating `traj_per_core` trajectories.  Returns (state-steps/s, sample string)."""
    import multiprocessing as mp
    from ozone_b200 import _native as nat
    d = nat.rhs_info(nat.RHS_IDS[rhs_name])['d']
    jobs = []
    for c in range(cores):
        case = synth_inputs(rhs_name, traj_per_core, nt, horizon, seed=100 + c)
        jobs.append((rhs_name, method, case))
    ctx = mp.get_context('fork')
    best = float('inf')
    with ctx.Pool(cores) as pool:
        pool.map(_oracle_chunk, jobs[:cores])          # warm-up (imports, page-in)
        for _ in range(repeats):
            t0 = time.perf_counter()
            pool.map(_oracle_chunk, jobs)
            best = min(best, time.perf_counter() - t0)
    units = cores * traj_per_core * d * (nt - 1)
    sample = (f'{cores} processes x {traj_per_core} trajectories, numpy oracle port (OPENBLAS_NUM_THREADS=1 each), '
              f'forward + adjoint, best of {repeats}')
    return units / best, sample, best


def run_reference(args, rank, world):
    """--impl reference: the reference algorithm on the host cores.  The reference itself cannot run
    (csdl_alpha is an un-vendored, uninstallable dependency), so this times the oracle port."""
    if rank != 0:
        return
    rhs, method, N, nt, horizon = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    per_core = args.cpu_traj_per_core or 1024
    vals = []
    for _ in range(max(args.warmup, 0)):
        cpu_port_throughput(rhs, method, nt, horizon, cores, max(per_core // 4, 8))
    t_all = 0.0
    for _ in range(args.steps):
        v, sample, dt = cpu_port_throughput(rhs, method, nt, horizon, cores, per_core)
        vals.append(v)
        t_all += dt
    value = float(np.mean(vals))
    line = {
        'impl': 'reference', 'metric': 'state-steps/sec (fwd+adjoint)', 'value': value, 'unit': 'state-steps/s',
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * t_all / args.steps,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': args.workload, 'note': 'each step is a bounded sample of the workload: '
                   f'{cores * per_core} of {N} trajectories, same nt/method'},
        'cpu_baseline': {'value': value, 'unit': 'state-steps/s', 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': 'state-steps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument('--cpu-traj-per-core', type=int, default=None,
                    help='CPU sample size per core (default 4096 for cpu_baseline, 1024 per step for --impl reference)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--newton-tol', type=float, default=1e-7)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else args.warmup

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if args.impl == 'reference':
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from ozone_b200 import engine, _native as nat
    from ozone_b200.glms import get_integration_method

    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: ozone_b200 has no CPU path')
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)

    rhs, method, N, nt, horizon = WORKLOADS[args.workload]
    fp32 = args.workload.endswith('fp32')
    dtype = torch.float32 if fp32 else torch.float64
    w = 4 if fp32 else 8
    n = nt - 1
    case = synth_inputs(rhs, N, nt, horizon, seed=rank)
    p_shared = case['p'].shape[-1] == 1
    spec = engine.SolveSpec(rhs_name=rhs, method=get_integration_method(method), N=N, n=n, dtype=dtype, h_shared=True,
                            p_shared=p_shared, th_shared=False, store_history=True, newton_tol=args.newton_tol)
    d = spec.d
    want = ('h', 'p', 'th') if rhs == 'lotka_volterra' else ('h', 'p')
    plan = engine.TmPlan(spec, device=dev, grads=want, cotangents=('final',))

    # host (pinned) inputs and results for the e2e leg
    def pin(a):
        return torch.tensor(np.ascontiguousarray(a), dtype=dtype).pin_memory()
    host = {'y0': pin(case['y0']), 'h': pin(case['h']), 'p': pin(case['p'][:, :, 0] if p_shared else case['p'])}
    if case['th'] is not None:
        host['th'] = pin(case['th'])
    for k, v in host.items():
        getattr(plan, k).copy_(v)
    plan.final_bar.zero_()
    plan.final_bar[-1].fill_(1.0) if rhs == 'van_der_pol' else plan.final_bar.fill_(1.0)
    host_out = {k: torch.empty_like(g, device='cpu').pin_memory() for k, g in plan.grad.items() if g is not None}
    host_obj = torch.empty(N, dtype=dtype).pin_memory()
    h2d = sum(host[k].numel() * w for k in ('h', 'p'))
    d2h = sum(v.numel() * w for v in host_out.values()) + host_obj.numel() * w

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2

    def flush_l2():
        flush.fill_(rank & 0xff)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allreduce_shared():
        if world > 1 and plan.grad['h'] is not None:
            dist.all_reduce(plan.grad['h'])

    ev = lambda: torch.cuda.Event(enable_timing=True)

    def device_step(times):
        flush_l2()
        e0, e1, e2 = ev(), ev(), ev()
        e0.record(); plan.forward(); e1.record(); plan.adjoint(); allreduce_shared(); e2.record()
        times.append((e0, e1, e2))

    def e2e_step(times):
        flush_l2()
        e0, e1 = ev(), ev()
        e0.record()
        plan.h.copy_(host['h'], non_blocking=True)
        plan.p.copy_(host['p'], non_blocking=True)
        plan.forward()
        plan.adjoint()
        allreduce_shared()
        host_obj.copy_(plan.final[-1], non_blocking=True)
        for k, v in host_out.items():
            v.copy_(plan.grad[k], non_blocking=True)
        e1.record()
        times.append((e0, e1))

    for _ in range(args.warmup):
        device_step([])
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    times = []
    barrier()
    t_wall0 = time.perf_counter()
    for _ in range(args.steps):
        device_step(times)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    fwd_ms = [a.elapsed_time(b) for a, b, _ in times]
    adj_ms = [b.elapsed_time(c) for _, b, c in times]
    step_ms = float(np.sum(fwd_ms) + np.sum(adj_ms))

    for _ in range(args.warmup):
        e2e_step([])
    barrier()
    etimes = []
    for _ in range(args.steps):
        e2e_step(etimes)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    e2e_ms = float(np.sum([a.elapsed_time(b) for a, b in etimes]))

    status_bad = int((plan.status != 0).sum().item())
    mean_newton = float(plan.iters.double().mean().item()) / n

    tot = torch.tensor([step_ms, e2e_ms, float(np.sum(fwd_ms)), float(np.sum(adj_ms))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    step_ms, e2e_ms, fwd_tot, adj_tot = tot.tolist()

    units_per_step = world * N * d * n
    value = units_per_step * args.steps / (step_ms * 1e-3)
    e2e_value = units_per_step * args.steps / (e2e_ms * 1e-3)

    if rank == 0:
        info = spec.info
        fwd_b, adj_b = algorithmic_bytes(info, N, n, w, p_shared, hist_cot=False)
        adj_avg_s = adj_tot / args.steps * 1e-3
        fwd_avg_s = fwd_tot / args.steps * 1e-3
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        except Exception:
            pass
        peak = float(peaks.get('hbm_gbs', 6650.0))
        achieved = adj_b / adj_avg_s / 1e9
        roof = {'bound': 'hbm', 'kernel': 'tm_adjoint_kernel', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                'frac': achieved / peak, 'traffic': None,
                'peak_source': 'MEASURED_PEAKS.json hbm_gbs (burst copy)' if peaks else 'fallback 6650 GB/s',
                'algorithmic_bytes_per_launch': adj_b, 'avg_launch_ms': adj_avg_s * 1e3,
                'forward_kernel': {'achieved': fwd_b / fwd_avg_s / 1e9, 'frac': fwd_b / fwd_avg_s / 1e9 / peak,
                                   'algorithmic_bytes_per_launch': fwd_b, 'avg_launch_ms': fwd_avg_s * 1e3},
                'note': 'implicit GL + Newton is FP64-ALU bound, not HBM bound (DESIGN.md); HBM fraction reported as the metric asks'
                        if not spec.method.explicit else 'explicit RK step is HBM/FP64 balanced'}
        cpu = None
        if not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            v, sample, _ = cpu_port_throughput(rhs, method, nt, horizon, cores, args.cpu_traj_per_core or 4096)
            cpu = {'value': v, 'unit': 'state-steps/s', 'cores': cores, 'kind': 'port', 'sample': sample}
        line = {
            'metric': 'state-steps/sec (fwd+adjoint)', 'value': value, 'unit': 'state-steps/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': step_ms / args.steps, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32' if fp32 else 'f64', 'data': 'synthetic',
            'config': {'workload': args.workload, 'rhs': rhs, 'method': method, 'approach': 'TimeMarching + adjoint',
                       'trajectories_per_gpu': N, 'num_times': nt, 'state_size': d, 'newton_tol': args.newton_tol,
                       'mean_newton_iters_per_step': mean_newton, 'l2': 'flushed (256 MiB write) between timed steps',
                       'parallelism': f'trajectory-batch x{world}', 'not_converged': status_bad},
            'e2e': {'value': e2e_value, 'unit': 'state-steps/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                    'ms_per_step': e2e_ms / args.steps},
            'gpu_launches': args.steps * (plan.launches_forward + plan.launches_adjoint),
            'roofline': roof, 'cpu_baseline': cpu, 'clocks': clocks,
            'wall_ms_timed_loop': t_wall * 1e3,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
