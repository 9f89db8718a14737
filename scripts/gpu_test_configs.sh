#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gputests.log
tail -n 5 gpurun_out/gputests.log
timeout 1200 python scripts/bench_configs.py "$@" > gpurun_out/configs.log 2>&1
echo "configs rc=$?"
