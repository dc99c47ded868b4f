#!/usr/bin/env python
"""Condense an .ncu-rep (via `ncu -i … --page raw --csv`) to one line per profiled launch."""
import csv, io, subprocess, sys
WANT = [('gpu__time_duration.sum', 'dur'), ('dram__bytes_read.sum', 'rd'), ('dram__bytes_write.sum', 'wr'),
        ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram%'),
        ('sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'fp64pipe%'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ%'), ('launch__registers_per_thread', 'regs'),
        ('smsp__inst_executed.sum', 'inst')]
for rep in sys.argv[1:]:
    txt = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print('##', rep)
    for r in rows[2:]:
        name = r[idx['Kernel Name']].split('(')[0][:60]
        parts = [f'{lab}={r[idx[m]]}{units[idx[m]]}' for m, lab in WANT if m in idx]
        print('  ', name, ' '.join(parts))
