#!/bin/bash
# GPU box: ONE compute-sanitizer tool per call (B200_PROFILING.md) over a small-shape subset of the GPU tests that goes through
# every kernel family: K1 (staged / direct / DMMA), K2, K3, the fused model kernels (staged, 128-bit lanes, warp per problem,
# DMMA, GD flavours) with their mbarrier / bulk-copy rings, and the retire-mode work lists.
# usage: gpu_sanitize.sh memcheck|racecheck|initcheck|synccheck   -> gpurun_out/sanitize_<tool>.log
set -u
TOOL=${1:-memcheck}
mkdir -p gpurun_out
LOG=gpurun_out/sanitize_${TOOL}.log
TESTS="tests/test_gpu_solve.py
tests/test_gpu_parity.py::test_normal_eq_matches_oracle
tests/test_gpu_parity.py::test_fused_normal_eq_matches_oracle
tests/test_gpu_parity.py::test_speculative_lm_schedule_is_bit_identical_to_two_pass
tests/test_gpu_multi_emg.py::test_multi_emg_tensor_core_fused_normal_equations_match_the_oracle
tests/test_gpu_retire.py::test_retire_freezes_each_problem_at_its_own_stop_iteration_synthetic
tests/test_gpu_retire.py::test_retire_warp_per_problem_kernels_skew_gauss"
# the plain run first: the sanitizer only sees a program that passes without it
python -m pytest $TESTS -m gpu -q -x > gpurun_out/sanitize_plain.log 2>&1 || { tail -20 gpurun_out/sanitize_plain.log; echo "plain run failed: not sanitized"; exit 1; }
EXTRA=""
[ "$TOOL" = racecheck ] && EXTRA="--racecheck-report all"
[ "$TOOL" = initcheck ] && EXTRA="--track-unused-memory no"
timeout 1500 compute-sanitizer --tool $TOOL $EXTRA --error-exitcode 99 --target-processes all \
  python -m pytest $TESTS -m gpu -q -x -p no:cacheprovider > $LOG 2>&1
echo "compute-sanitizer --tool $TOOL exit code: $?" | tee -a $LOG
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" $LOG | tail -8
