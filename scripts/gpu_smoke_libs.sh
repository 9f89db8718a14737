#!/bin/bash
# GPU box: smoke() under each of the given libraries (NLLS_LIB).  usage: gpu_smoke_libs.sh "<lib.so ...>"
mkdir -p gpurun_out
for l in $1; do
  t=$(basename $l .so)
  NLLS_LIB=$PWD/$l timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$t.log 2>&1
  echo "== $t rc=$?"; tail -2 gpurun_out/smoke_$t.log
done
