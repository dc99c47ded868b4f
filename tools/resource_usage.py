#!/usr/bin/env python
"""Summarise registers / stack (spills) / shared memory of every kernel in liboz_b200.so."""
import re, subprocess, sys
so = sys.argv[1] if len(sys.argv) > 1 else 'ozone_b200/liboz_b200.so'
txt = subprocess.run(['cuobjdump', '--dump-resource-usage', so], capture_output=True, text=True).stdout
rows = []
for m in re.finditer(r'Function (\S+):\s*\n\s*(REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)[^\n]*)', txt):
    name = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip()
    name = re.sub(r'\(.*', '', name).replace('void oz::', '')
    rows.append((int(m.group(3)), int(m.group(4)), int(m.group(5)), name))
pat = sys.argv[2] if len(sys.argv) > 2 else ''
for r in sorted(rows, key=lambda r: (-r[1], -r[0])):
    if pat in r[3]:
        print(f'REG {r[0]:4d} STACK {r[1]:6d} SMEM {r[2]:6d}  {r[3]}')
