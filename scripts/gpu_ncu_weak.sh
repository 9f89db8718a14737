#!/bin/bash
# GPU box: ncu full captures (one launch each) of the kernels the config sweep showed furthest from the HBM roofline.
# The .ncu-rep files are reduced to raw + source CSV summaries on the box (gpurun_out/ is capped at 64 MiB).
set -u
mkdir -p gpurun_out
cap() {  # name case kernel-regex skip
  python scripts/bench_configs.py $2 > gpurun_out/plain_$1.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$3 -s $4 -c 1 -o /tmp/prof_$1 -f python scripts/bench_configs.py $2 > gpurun_out/ncu_$1.log 2>&1
  echo "$1 rc=$?"
  ncu -i /tmp/prof_$1.ncu-rep --page raw --csv > /tmp/raw_$1.csv 2>/dev/null
  ncu -i /tmp/prof_$1.ncu-rep --page source --csv > /tmp/src_$1.csv 2>/dev/null
  { python scripts/ncu_raw_summary.py /tmp/raw_$1.csv; python scripts/ncu_source_summary.py /tmp/src_$1.csv 30; } > gpurun_out/sum_$1.txt 2>&1
}
for spec in "$@"; do
  IFS=: read -r name cs rx skip <<< "$spec"
  cap "$name" "$cs" "$rx" "$skip"
done
