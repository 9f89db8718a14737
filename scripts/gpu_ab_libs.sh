#!/bin/bash
# GPU box: A/B of whole libraries (NLLS_LIB) over cases of scripts/bench_configs.py.  usage: gpu_ab_libs.sh "<lib.so ...>" <case> ...
set -u
mkdir -p gpurun_out
LIBS=$1; shift
for l in $LIBS; do
  t=$(basename $l .so)
  NLLS_LIB=$PWD/$l timeout 600 python scripts/bench_configs.py "$@" > gpurun_out/ab_$t.jsonl 2>gpurun_out/ab_$t.err
  echo "== $t rc=$?"
  python scripts/show_configs.py gpurun_out/ab_$t.jsonl
done
