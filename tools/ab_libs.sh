#!/bin/bash
# A/B of library builds on one workload:  tools/ab_libs.sh <workload> lib1.so lib2.so ...   (each run twice, interleaved)
wl=$1; shift
for i in 1 2; do
for lib in "$@"; do
  OZONE_B200_LIB=$PWD/$lib python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$lib', '%.4g' % d['value'], r['kernel'], '%.4f' % r['avg_launch_ms'], r['other_kernel']['kernel'], '%.4f' % r['other_kernel']['avg_launch_ms'])"
done; done
