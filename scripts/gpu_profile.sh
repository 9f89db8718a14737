#!/bin/bash
# Runs on the GPU box (under gpurun): full GPU tests, both bench paths, ncu launch lists and
# full captures of the two dominant kernels.  Outputs land in gpurun_out/.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gputests.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_fused.log 2>&1; echo "rc=$?" >> gpurun_out/bench_fused.log
python bench.py --steps 3 --warmup 3 --path tensor --no-cpu-baseline > gpurun_out/bench_tensor.log 2>&1; echo "rc=$?" >> gpurun_out/bench_tensor.log
A="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --path tensor"
$A > gpurun_out/plainA.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_fused.csv $A > gpurun_out/ncuA.log 2>&1
$A > gpurun_out/plainA2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_model_normal_eq -s 30 -c 3 -o gpurun_out/prof_fused $A > gpurun_out/ncuA2.log 2>&1
$B > gpurun_out/plainB.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_tensor.csv $B > gpurun_out/ncuB.log 2>&1
$B > gpurun_out/plainB2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_normal_eq -s 30 -c 3 -o gpurun_out/prof_tensor $B > gpurun_out/ncuB2.log 2>&1
ls -la gpurun_out
