#!/bin/bash
# GPU box: ncu launch list + full capture of the fused normal-equation kernel in the bench workload.
# usage: gpu_ncu_fused.sh <tag>     (outputs gpurun_out/<tag>_*)
set -u
TAG=${1:-r02}
mkdir -p gpurun_out
A="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$A > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $A > gpurun_out/${TAG}_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_model_normal_eq -s 30 -c 2 -o gpurun_out/${TAG}_prof_fused $A > gpurun_out/${TAG}_ncu_full.log 2>&1
cut -c1-600 gpurun_out/${TAG}_plain.log
ls -la gpurun_out/
