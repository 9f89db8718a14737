#!/bin/bash
# GPU box: tests, then ncu full capture of the fused normal-equation kernel in the bench workload.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gputests.log
tail -n 3 gpurun_out/gputests.log
A="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$A > gpurun_out/plainA.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_model_normal_eq -s 30 -c 2 -o gpurun_out/prof_fused $A > gpurun_out/ncuA2.log 2>&1
cat gpurun_out/plainA.log | cut -c1-400
