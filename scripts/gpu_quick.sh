#!/bin/bash
# Runs on the GPU box: GPU tests + both benches (no profiler).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/gputests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gputests.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_fused.log 2>&1; echo "rc=$?" >> gpurun_out/bench_fused.log
tail -n 4 gpurun_out/gputests.log; cat gpurun_out/bench_fused.log
