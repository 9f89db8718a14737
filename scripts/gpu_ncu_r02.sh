#!/bin/bash
# GPU box: ncu --set full captures of the dominant kernels of every single-GPU config (one gpurun call; each command has
# exited 0 without ncu right before).  usage: gpu_ncu_r02.sh <tag>   -> gpurun_out/<tag>_*.ncu-rep
set -u
TAG=${1:-r02n}
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
# a report is ~28 MB and gpurun_out/ travels back only below 64 MiB: summarise on the box, keep the raw metric table
digest() {   # digest <name> [keep-source-csv]
  local R=gpurun_out/${TAG}_$1
  [ -f $R.ncu-rep ] || { echo "no report for $1"; return; }
  ncu -i $R.ncu-rep --page raw --csv > $R.raw.csv 2>/dev/null
  ncu -i $R.ncu-rep --page source --csv > $R.src.csv 2>/dev/null
  { python scripts/ncu_raw_summary.py $R.raw.csv; python scripts/ncu_source_summary.py $R.src.csv 40; } > $R.summary.txt 2>&1
  [ "${2:-}" = keep ] || rm -f $R.src.csv
  rm -f $R.ncu-rep
}
A="python bench.py --steps 1 --warmup 2 --no-cpu-baseline"
$A > gpurun_out/${TAG}_plain_bench.log 2>&1 && {
  ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 300 --csv --log-file gpurun_out/${TAG}_launches.csv $A > /dev/null 2>&1
  $NCU -k regex:k_model_normal_eq_staged -s 30 -c 2 -o gpurun_out/${TAG}_fused_gn $A > gpurun_out/${TAG}_ncu1.log 2>&1; digest fused_gn keep
  $NCU -k regex:k_normal_eq_staged -s 4 -c 2 -o gpurun_out/${TAG}_k1 $A > gpurun_out/${TAG}_ncu2.log 2>&1; digest k1
}
Bc="python scripts/bench_configs.py cfg2-gd"
$Bc > gpurun_out/${TAG}_plain_gd.log 2>&1 && $NCU -k regex:k_model_normal_eq_gda -s 20 -c 2 -o gpurun_out/${TAG}_fused_gd $Bc > gpurun_out/${TAG}_ncu3.log 2>&1; digest fused_gd
Cc="python scripts/bench_configs.py cfg3-lm"
$Cc > gpurun_out/${TAG}_plain_c3lm.log 2>&1 && $NCU -k regex:k_model_wpp -s 20 -c 2 -o gpurun_out/${TAG}_wpp_neq $Cc > gpurun_out/${TAG}_ncu4.log 2>&1; digest wpp_neq
Dc="python scripts/bench_configs.py cfg3-gd"
$Dc > gpurun_out/${TAG}_plain_c3gd.log 2>&1 && $NCU -k regex:k_model_wpp -s 20 -c 2 -o gpurun_out/${TAG}_wpp_gda $Dc > gpurun_out/${TAG}_ncu5.log 2>&1; digest wpp_gda
Ec="python scripts/bench_configs.py cfg2-gn-f32"
$Ec > gpurun_out/${TAG}_plain_f32.log 2>&1 && $NCU -k regex:k_model_normal_eq_vec -s 20 -c 2 -o gpurun_out/${TAG}_vec_f32 $Ec > gpurun_out/${TAG}_ncu6.log 2>&1; digest vec_f32
Fc="python scripts/bench_configs.py cfg4-gn"
$Fc > gpurun_out/${TAG}_plain_c4.log 2>&1 && $NCU -k regex:k_model_normal_eq_mma -s 6 -c 2 -o gpurun_out/${TAG}_mma $Fc > gpurun_out/${TAG}_ncu7.log 2>&1; digest mma
ls -la gpurun_out/ | grep ${TAG}
