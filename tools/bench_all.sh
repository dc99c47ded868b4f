#!/bin/bash
# Every bench line of a round on one GPU:  tools/bench_all.sh <tag>   ->  gpurun_out/<tag>_{default.json,reference.json,others.jsonl}
tag=${1:-r2}
set -x
python bench.py --steps 20 --warmup 5 > gpurun_out/${tag}_default.json 2> gpurun_out/${tag}_default.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_reference.json 2> gpurun_out/${tag}_reference.err
rm -f gpurun_out/${tag}_others.jsonl
for w in van_der_pol_gl4_tm_adjoint_N65536_nt40 van_der_pol_gl2_tm_adjoint_N65536_nt40 lotka_volterra_rk4_tm_adjoint_N1M_nt101 lotka_volterra_rk4_tm_adjoint_N1M_nt101_fp32 lotka_volterra_rk4_picard_adjoint_N1M_nt101 lotka_volterra_rk4_picard_jacobi_adjoint_N1M_nt101 lotka_volterra_rk4_picard_adjoint_N1M_nt101_fp32 diffusion_reaction_rk4_checkpoints_nx1M_n1024 diffusion_reaction_rk4_checkpoints_nx1M_n1024_fp32 diffusion_reaction_rk4_checkpoints_nx1M_n1024_traced diffusion_reaction_rk4_tm_history_nx1M_n1024 diffusion_reaction_rk4_checkpoints_nx1M_n1024_sharded lunar_ascent_rk4_collocation_N16384_nt101 lunar_ascent_gl6_collocation_N16384_nt101 lunar_ascent_rk4_optimizer_eval_N16384_nt101; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline >> gpurun_out/${tag}_others.jsonl 2>> gpurun_out/${tag}_others.err
done
tail -c 400 gpurun_out/${tag}_default.json
