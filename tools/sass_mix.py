"""Instruction mix of address ranges in a `cuobjdump -sass` listing (used to count what one loop iteration issues).

    python tools/sass_mix.py listing.sass 0x1ac0-0x2200 0x2460-0x28f0 ...
"""
import collections
import re
import sys


def main():
    path, ranges = sys.argv[1], [tuple(int(x, 16) for x in r.split('-')) for r in sys.argv[2:]]
    mix = collections.Counter()
    for line in open(path):
        m = re.match(r'\s+/\*([0-9a-f]{4,5})\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)', line)
        if m and any(lo <= int(m.group(1), 16) < hi for lo, hi in ranges):
            mix[m.group(2).split('.')[0]] += 1
    fp64 = sum(mix[k] for k in ('DFMA', 'DADD', 'DMUL'))
    print(f"fp64 {fp64} (DFMA {mix['DFMA']} DADD {mix['DADD']} DMUL {mix['DMUL']}), "
          f"fp32 {sum(mix[k] for k in ('FFMA', 'FADD', 'FMUL'))}, other {sum(mix.values()) - fp64}, total {sum(mix.values())}")
    print(' '.join(f'{k}:{v}' for k, v in mix.most_common()))


if __name__ == '__main__':
    main()
