#!/bin/bash
# GPU box: sweep NLLS_FUSED_VARIANT over cases of scripts/bench_configs.py.  usage: gpu_sweep.sh "<variants>" <case> ...
set -u
mkdir -p gpurun_out
VARS=$1; shift
for v in $VARS; do
  NLLS_FUSED_VARIANT=$v timeout 600 python scripts/bench_configs.py "$@" > gpurun_out/sweep_v$v.jsonl 2>gpurun_out/sweep_v$v.err
  echo "== variant $v rc=$?"
  python scripts/show_configs.py gpurun_out/sweep_v$v.jsonl
done
