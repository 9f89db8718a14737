#!/bin/bash
# GPU box: sweep NLLS_FUSED_VARIANT over given bench_configs cases (no profiler).
# usage: gpu_variants.sh <tag> "<variants>" "<cases>"
set -u
TAG=${1:-r02}; VARS=${2:-"0 1 2 3"}; CASES=${3:-"cfg2-gn"}
mkdir -p gpurun_out
OUT=gpurun_out/${TAG}_variants.jsonl
: > $OUT
for v in $VARS; do
  echo "{\"variant\": $v}" >> $OUT
  NLLS_FUSED_VARIANT=$v timeout 300 python scripts/bench_configs.py $CASES >> $OUT 2>> gpurun_out/${TAG}_variants.err
done
python scripts/show_configs.py $OUT
