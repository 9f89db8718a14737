#!/bin/bash
# Runs on the GPU box (under gpurun): full GPU tests, both bench paths, ncu launch lists and full captures of the
# dominant kernels.  The .ncu-rep files stay on the box (/tmp): gpurun_out/ is capped at 64 MiB, so they are reduced
# here to launch summaries + raw-metric / source-page summaries (scripts/ncu_*.py).
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gputests.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_fused.log 2>&1; echo "rc=$?" >> gpurun_out/bench_fused.log
python bench.py --steps 3 --warmup 3 --path tensor --no-cpu-baseline > gpurun_out/bench_tensor.log 2>&1; echo "rc=$?" >> gpurun_out/bench_tensor.log
A="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --path tensor"
summarise() {  # name
  ncu -i /tmp/prof_$1.ncu-rep --page raw --csv > /tmp/raw_$1.csv 2>/dev/null
  ncu -i /tmp/prof_$1.ncu-rep --page source --csv > /tmp/src_$1.csv 2>/dev/null
  { python scripts/ncu_raw_summary.py /tmp/raw_$1.csv; python scripts/ncu_source_summary.py /tmp/src_$1.csv 30; } > gpurun_out/sum_$1.txt 2>&1
}
$A > gpurun_out/plainA.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file /tmp/launches_fused.csv $A > gpurun_out/ncuA.log 2>&1
python scripts/ncu_launch_summary.py /tmp/launches_fused.csv > gpurun_out/launches_fused.txt 2>&1
$A > gpurun_out/plainA2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_model_normal_eq -s 30 -c 2 -f -o /tmp/prof_fused $A > gpurun_out/ncuA2.log 2>&1
summarise fused
ncu --set full --clock-control none --import-source on -k regex:k_model_normal_eq -s 130 -c 1 -f -o /tmp/prof_lmtrial $A > gpurun_out/ncuA3.log 2>&1
summarise lmtrial
$B > gpurun_out/plainB.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file /tmp/launches_tensor.csv $B > gpurun_out/ncuB.log 2>&1
python scripts/ncu_launch_summary.py /tmp/launches_tensor.csv > gpurun_out/launches_tensor.txt 2>&1
$B > gpurun_out/plainB2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_normal_eq -s 30 -c 2 -f -o /tmp/prof_tensor $B > gpurun_out/ncuB2.log 2>&1
summarise tensor
ls -la gpurun_out
