#!/bin/bash
# GPU box: A/B of whole libraries on the cfg-2 cases, then the parity tests that see the fp64 exp on the in-tree build, then one
# ncu --set full capture of the in-tree build's fp64 fused kernel with the per-SASS-line table kept.  usage: gpu_ab_exp.sh <tag> "<libs>"
set -u
TAG=$1; LIBS=$2
mkdir -p gpurun_out
for rep in 1 2; do
for l in $LIBS; do
  t=$(basename $l .so)
  NLLS_LIB=$PWD/$l timeout 600 python scripts/bench_configs.py cfg2-gn cfg2-lm cfg2-gd cfg4-gn > gpurun_out/${TAG}_ab_${t}_$rep.jsonl 2>gpurun_out/${TAG}_ab_${t}_$rep.err
  echo "== $t rep $rep rc=$?"
  python scripts/show_configs.py gpurun_out/${TAG}_ab_${t}_$rep.jsonl | grep -v "nlls_lm_solve\|update\|iter_end"
done
done
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_single.py tests/test_gpu_retire.py -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest: $?"; tail -4 gpurun_out/${TAG}_pytest.log
A="python scripts/bench_configs.py cfg2-gn"
R=gpurun_out/${TAG}_fused_gn
ncu --set full --clock-control none --import-source on -k regex:k_model_normal_eq_staged -s 20 -c 1 -o $R $A > gpurun_out/${TAG}_ncu.log 2>&1
ncu -i $R.ncu-rep --page raw --csv > $R.raw.csv 2>/dev/null
ncu -i $R.ncu-rep --page source --csv > $R.src.csv 2>/dev/null
{ python scripts/ncu_raw_summary.py $R.raw.csv; python scripts/ncu_source_summary.py $R.src.csv 40; } > $R.summary.txt 2>&1
rm -f $R.ncu-rep
ls -la gpurun_out | grep $TAG
