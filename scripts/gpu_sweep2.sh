#!/bin/bash
# GPU box: every single-GPU config case (per-kernel times) + NLLS_FUSED_VARIANT sweep of the cfg-2 kernels. No profiler.
# usage: gpu_sweep2.sh <tag>
set -u
TAG=${1:-r02}
mkdir -p gpurun_out
OUT=gpurun_out/${TAG}_configs.jsonl
: > $OUT
timeout 900 python scripts/bench_configs.py >> $OUT 2> gpurun_out/${TAG}_configs.err
for v in 1 2 3 4 5 6 7; do
  echo "{\"variant\": $v}" >> $OUT
  NLLS_FUSED_VARIANT=$v timeout 300 python scripts/bench_configs.py cfg2-gn cfg2-gn-f32 >> $OUT 2>> gpurun_out/${TAG}_configs.err
done
python scripts/show_configs.py $OUT 2>/dev/null || cut -c1-400 $OUT
