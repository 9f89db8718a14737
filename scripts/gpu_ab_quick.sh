#!/bin/bash
# GPU box: alternating A/B of whole libraries on bench_configs cases (kernel averages only).  usage: gpu_ab_quick.sh <tag> "<libs>" "<cases>" [reps]
set -u
TAG=$1; LIBS=$2; CASES=$3; REPS=${4:-3}
mkdir -p gpurun_out
for rep in $(seq 1 $REPS); do
for l in $LIBS; do
  t=$(basename $l .so)
  NLLS_LIB=$PWD/$l timeout 600 python scripts/bench_configs.py $CASES > gpurun_out/${TAG}_ab_${t}_$rep.jsonl 2>gpurun_out/${TAG}_ab_${t}_$rep.err
  echo "== $t rep $rep rc=$?"
  python scripts/show_configs.py gpurun_out/${TAG}_ab_${t}_$rep.jsonl | grep "model" 
done
done
