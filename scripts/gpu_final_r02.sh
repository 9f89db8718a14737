#!/bin/bash
# GPU box, one gpurun call: the no-allocation entry point's test alone (fresh process), the whole GPU suite, the 1-GPU bench
# record, then the ncu launch list and --set full captures of the kernels bench.py reports a roofline for (digested on the box:
# gpurun_out/ travels back only below 64 MiB).  usage: gpu_final_r02.sh <tag>
set -u
TAG=${1:-r02q}
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/${TAG}_zz_alone.log 2>&1; echo "zz alone: $?"
tail -3 gpurun_out/${TAG}_zz_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest -m gpu: $?"
tail -5 gpurun_out/${TAG}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke: $?"
python bench.py > gpurun_out/${TAG}_bench_1gpu.json 2> gpurun_out/${TAG}_bench_1gpu.err; echo "bench: $?"
tail -c 1500 gpurun_out/${TAG}_bench_1gpu.json
NCU="ncu --set full --clock-control none --import-source on"
digest() {   # digest <name>
  local R=gpurun_out/${TAG}_$1
  [ -f $R.ncu-rep ] || { echo "no report for $1"; return; }
  ncu -i $R.ncu-rep --page raw --csv > $R.raw.csv 2>/dev/null
  ncu -i $R.ncu-rep --page source --csv > $R.src.csv 2>/dev/null
  { python scripts/ncu_raw_summary.py $R.raw.csv; python scripts/ncu_source_summary.py $R.src.csv 40; } > $R.summary.txt 2>&1
  rm -f $R.src.csv $R.ncu-rep
}
A="python bench.py --steps 1 --warmup 2 --no-cpu-baseline --no-cfg5"
$A > gpurun_out/${TAG}_plain_bench.log 2>&1 && {
  ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 300 --csv --log-file gpurun_out/${TAG}_launches.csv $A > /dev/null 2>&1
  $NCU -k regex:k_model_normal_eq_staged -s 30 -c 2 -o gpurun_out/${TAG}_fused_gn $A > gpurun_out/${TAG}_ncu1.log 2>&1; digest fused_gn
  $NCU -k regex:k_normal_eq_staged -s 4 -c 2 -o gpurun_out/${TAG}_k1 $A > gpurun_out/${TAG}_ncu2.log 2>&1; digest k1
}
L="python scripts/bench_configs.py cfg2-lm"
$L > gpurun_out/${TAG}_plain_lm.log 2>&1 && $NCU -k regex:k_model_normal_eq_staged -s 30 -c 2 -o gpurun_out/${TAG}_fused_lm $L > gpurun_out/${TAG}_ncu3.log 2>&1; digest fused_lm
Ec="python scripts/bench_configs.py cfg2-gn-f32"
$Ec > gpurun_out/${TAG}_plain_f32.log 2>&1 && $NCU -k regex:k_model_normal_eq_vec -s 20 -c 2 -o gpurun_out/${TAG}_vec_f32 $Ec > gpurun_out/${TAG}_ncu4.log 2>&1; digest vec_f32
Bc="python scripts/bench_configs.py cfg2-gd"
$Bc > gpurun_out/${TAG}_plain_gd.log 2>&1 && $NCU -k regex:k_model_normal_eq_gda -s 20 -c 2 -o gpurun_out/${TAG}_fused_gd $Bc > gpurun_out/${TAG}_ncu5.log 2>&1; digest fused_gd
ls -la gpurun_out/ | grep ${TAG}
