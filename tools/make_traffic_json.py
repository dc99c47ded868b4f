#!/usr/bin/env python
"""Regenerate profiles/traffic.json from `ncu --set full` reports of the SHIPPED build (run on the GPU box, where the
reports are; only this JSON and the text summaries travel back).

    python tools/make_traffic_json.py <tag> <dir with vdp.ncu-rep lv.ncu-rep ...> > traffic.json

Per workload and kernel: DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum, averaged over the
captured launches) and, for fp64 kernels, the FP64 flops executed per launch from the source page
(2 x DFMA + DADD + DMUL predicated-on thread instructions).  bench.py copies them into `roofline`."""
import collections
import csv
import io
import json
import os
import re
import subprocess
import sys

FAMILIES = {   # report name -> (workload, {kernel-name regex: key in traffic.json})
    'vdp': ('van_der_pol_gl6_tm_adjoint_N65536_nt40', {'tm_forward_kernel': 'tm_forward_kernel', 'tm_adjoint_kernel': 'tm_adjoint_kernel'}),
    'lv': ('lotka_volterra_rk4_tm_adjoint_N1M_nt101', {'tm_forward_kernel': 'tm_forward_kernel', 'tm_adjoint_kernel': 'tm_adjoint_kernel'}),
    'lv32': ('lotka_volterra_rk4_tm_adjoint_N1M_nt101_fp32', {'tm_forward_kernel': 'tm_forward_kernel', 'tm_adjoint_kernel': 'tm_adjoint_kernel'}),
    'pde': ('diffusion_reaction_rk4_checkpoints_nx1M_n1024', {'pde_step_kernel': 'pde_step_kernel', 'pde_adjoint_step_kernel': 'pde_adjoint_step_kernel'}),
    'pde32': ('diffusion_reaction_rk4_checkpoints_nx1M_n1024_fp32', {'pde_step_kernel': 'pde_step_kernel', 'pde_adjoint_step_kernel': 'pde_adjoint_step_kernel'}),
    'pdetr': ('diffusion_reaction_rk4_checkpoints_nx1M_n1024_traced', {'pde_step_kernel': 'pde_step_kernel', 'pde_adjoint_step_kernel': 'pde_adjoint_step_kernel'}),
    'colloc': ('lunar_ascent_rk4_collocation_N16384_nt101', {'colloc_residual_kernel': 'colloc_residual_kernel', 'colloc_vjp_kernel': 'colloc_vjp_kernel'}),
    'coljac': ('lunar_ascent_rk4_optimizer_eval_N16384_nt101', {'colloc_residual_kernel': 'colloc_residual_kernel', 'rhs_jac_kernel': 'rhs_jac_kernel'}),
    'picard': ('lotka_volterra_rk4_picard_adjoint_N1M_nt101', {'picard_kernel': 'picard_kernel'}),
}


def fp64_flops(rep, kernel):
    txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', f'regex:{kernel}'],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    launches, flops, hdr = 0, 0, None
    for r in rows:
        if r and r[0] == 'Address':
            hdr = {h: i for i, h in enumerate(r)}
            launches += 1
            continue
        if hdr is None or len(r) < len(hdr):
            continue
        m = re.match(r'\s*(?:@!?U?P\d\s+)?([A-Z0-9_.]+)', r[hdr['Source']])
        n = r[hdr['Predicated-On Thread Instructions Executed']]
        if not m or not n.isdigit():
            continue
        op = m.group(1).split('.')[0]
        if op == 'DFMA':
            flops += 2 * int(n)
        elif op in ('DADD', 'DMUL'):
            flops += int(n)
    return flops // max(launches, 1)


def main():
    tag, d = sys.argv[1], sys.argv[2]
    out = {'_comment': f'DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) and executed FP64 flops per launch from the '
                       f'`ncu --set full` captures of build {tag} (tools/profile_all.sh -> tools/make_traffic_json.py; text summaries in '
                       f'profiles/{tag}_ncu_summary.txt); bench.py copies them into roofline.traffic / roofline.flops_per_launch.'}
    for fam, (workload, kernels) in FAMILIES.items():
        rep = os.path.join(d, fam + '.ncu-rep')
        if not os.path.exists(rep):
            continue
        entry = {'source': f'profiles/{tag}_ncu_summary.txt ({fam})'}
        flops = {}
        # the raw page has a units row under the header
        txt = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
        rws = list(csv.reader(io.StringIO(txt)))
        hdr, units = rws[0], rws[1]
        idx = {h: i for i, h in enumerate(hdr)}
        scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
        for pat, key in kernels.items():
            tot, cnt = 0.0, 0
            for r in rws[2:]:
                if not re.search(r'\b' + pat + r'\b', r[idx['Kernel Name']]):
                    continue
                for m in ('dram__bytes_read.sum', 'dram__bytes_write.sum'):
                    tot += float(r[idx[m]]) * scale.get(units[idx[m]], 1.0)
                cnt += 1
            if cnt:
                entry[key] = int(tot / cnt)
                if not workload.endswith('fp32'):
                    fl = fp64_flops(rep, pat)
                    if fl:
                        flops[key] = fl
        if flops:
            flops['source'] = f'source page of the same report: 2 x DFMA + DADD + DMUL predicated-on thread instructions per launch'
            entry['fp64_flops_per_launch'] = flops
        out[workload] = entry
    json.dump(out, sys.stdout, indent=2)
    print()


if __name__ == '__main__':
    main()
