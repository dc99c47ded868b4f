"""Executed-instruction mix of one kernel from an .ncu-rep captured with --import-source on.

    python tools/ncu_dynamic_mix.py report.ncu-rep regex:tm_forward

Static SASS listings over-count code on paths that never run (pivot swaps, fallbacks); this weights every
instruction by how often it was issued and shows where the stall samples fall.
"""
import collections
import csv
import io
import re
import subprocess
import sys


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', kernel],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    start = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
    hdr = rows[start]
    idx = {h: i for i, h in enumerate(hdr)}
    mix, samples, total = collections.Counter(), collections.Counter(), 0
    for r in rows[start + 1:]:
        if len(r) < len(hdr) or r[0] == 'Address' or not r[idx['Instructions Executed']].isdigit():
            continue
        m = re.match(r'\s*(?:@!?U?P\d\s+)?([A-Z0-9_.]+)', r[idx['Source']])
        if not m:
            continue
        op = m.group(1)
        key = 'IMAD.MOV' if op.startswith('IMAD.MOV') else op.split('.')[0]
        n = int(r[idx['Instructions Executed']])
        mix[key] += n
        total += n
        samples[key] += int(r[idx['# Samples']] or 0)
    print(f'{rows[0][1][:90]}: {total} warp instructions executed')
    s_tot = sum(samples.values()) or 1
    for k, v in mix.most_common(24):
        print(f'  {k:10s} {v:12d} {100 * v / total:5.1f} %   stall samples {100 * samples[k] / s_tot:5.1f} %')


if __name__ == '__main__':
    main()
