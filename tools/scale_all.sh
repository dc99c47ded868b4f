#!/bin/bash
# On an 8-GPU box: the default workload (trajectory batches, weak scaling) and the spatially sharded grid (strong scaling)
# at 1/2/4/8 GPUs, launched as the driver does.   tools/scale_all.sh <tag>  ->  gpurun_out/scale_<tag>/
out=gpurun_out/scale_$1; mkdir -p $out
W=diffusion_reaction_rk4_checkpoints_nx1M_n1024_sharded
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --no-other-workloads > $out/bench_n1.json 2> $out/bench_n1.err
python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline --workload $W > $out/shard_n1.json 2> $out/shard_n1.err
for n in 2 4 8; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n))"
  $TR bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline --no-other-workloads > $out/bench_n$n.json 2> $out/bench_n$n.err
  $TR bench.py --gpus $n --steps 5 --warmup 3 --no-cpu-baseline --workload $W > $out/shard_n$n.json 2> $out/shard_n$n.err
done
python - $out <<'PY'
import json, sys, glob, os
out = sys.argv[1]
for kind in ('bench', 'shard'):
    for n in (1, 2, 4, 8):
        try:
            d = json.loads(open(f'{out}/{kind}_n{n}.json').read().strip().splitlines()[-1])
            r = d['roofline']['hbm']
            print(kind, n, '%.4g' % d['value'], d['unit'], 'fwd %.4f adj %.4f ms' % (r['forward']['avg_launch_ms'], r['reverse']['avg_launch_ms']),
                  'e2e %.4g' % d['e2e']['value'], d['scaling'])
        except Exception as e:
            print(kind, n, 'failed', repr(e)[:100])
PY
