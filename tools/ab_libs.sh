for i in 1 2; do
for lib in build/liboz_old.so ozone_b200/liboz_b200.so; do
  OZONE_B200_LIB=$PWD/$lib python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$lib', d['value'], r['kernel'], r['avg_launch_ms'], r['other_kernel']['kernel'], r['other_kernel']['avg_launch_ms'])"
done; done
