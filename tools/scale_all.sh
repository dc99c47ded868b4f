#!/bin/bash
# On an 8-GPU box: default workload at 1/2/4/8 GPUs launched as the driver does, then the spatially sharded grid at 2/4/8.
# Also the same grid four times larger (the regime where a slab still fills a GPU).
out=gpurun_out/scale_$1; mkdir -p $out
python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > $out/bench_n1.json 2> $out/bench_n1.err
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 10 --warmup 3 --no-cpu-baseline > $out/bench_n$n.json 2> $out/bench_n$n.err
done
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) \
      tools/check_pde_shard.py > $out/pde_n$n.txt 2> $out/pde_n$n.err
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) \
      tools/check_pde_shard.py 4000000 256 16 > $out/pde4M_n$n.txt 2> $out/pde4M_n$n.err
done
cat $out/bench_n*.json | cut -c1-160; cat $out/pde*_n*.txt
