out=gpurun_out/scale_r1k; mkdir -p $out
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) tools/check_pde_shard.py > $out/pde_n$n.txt 2> $out/pde_n$n.err
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) tools/check_pde_shard.py 4000000 256 16 > $out/pde4M_n$n.txt 2> $out/pde4M_n$n.err
done
cat $out/pde*_n*.txt | grep -v NCCL
