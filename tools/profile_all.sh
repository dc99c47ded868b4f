#!/bin/bash
# ncu --set full for every kernel family at its bench size; summaries are made on the GPU box so that only text
# (plus the reports named in KEEP) comes back.   usage: tools/profile_all.sh <tag> [families...]
tag=$1; shift
fams=${@:-vdp lv lv32 pde pde32 colloc coljac picard}
KEEP=${KEEP:-vdp}
out=gpurun_out/prof_$tag
mkdir -p $out
for f in $fams; do
  python tools/ncu_targets.py $f > $out/plain_$f.log 2>&1 || { echo "plain run of $f failed"; continue; }
  ncu --set full --clock-control none --import-source on -k regex:"tm_|pde_.*step|colloc_|picard_|rhs_jac" -s 4 -c 4 -o $out/$f -f python tools/ncu_targets.py $f > $out/ncu_$f.log 2>&1 || echo "ncu $f failed"
  python tools/ncu_summary.py $out/$f.ncu-rep >> $out/summary.txt
  for k in tm_forward tm_adjoint pde_step pde_adjoint colloc_residual colloc_vjp picard_ rhs_jac; do
    python tools/ncu_dynamic_mix.py $out/$f.ncu-rep regex:$k >> $out/dynamic_mix.txt 2>/dev/null
  done
done
python tools/make_traffic_json.py $tag $out > $out/traffic.json
for f in $fams; do case " $KEEP " in *" $f "*) ;; *) rm -f $out/$f.ncu-rep;; esac; done
cat $out/summary.txt
