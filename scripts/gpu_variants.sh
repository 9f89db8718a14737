#!/bin/bash
# GPU box: sweep NLLS_FUSED_VARIANT over the bench workload (no profiler).
set -u
mkdir -p gpurun_out
for v in 0 1 2 3; do
  NLLS_FUSED_VARIANT=$v timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/variant_$v.log 2>&1
  echo "variant $v rc=$?"
  python - <<PY
import json
try:
    l=[x for x in open('gpurun_out/variant_$v.log') if x.startswith('{')][-1]
    d=json.loads(l); print('  value', round(d['value']), 'ms/launch', round(d['roofline']['avg_launch_ms'],4), 'frac', round(d['roofline']['frac'],4))
except Exception as e:
    print('  failed', e)
PY
done
