#!/bin/bash
# compute-sanitizer over the tests that exercise the code whose speed comes from removed bounds tests (padded fp32 map,
# biased base pointer, mantissa-floor addressing), shared-memory queues (k_agents, split kernels) and the new paths.
#   tools/sanitize.sh memcheck|racecheck     (ONE tool per gpurun call, see B200_PROFILING.md)
# Output: gpurun_out/sanitize_<tool>.log (+ a summary line per test); copy the summaries to profiles/.
set -u
TOOL=${1:-memcheck}
mkdir -p gpurun_out
OUT=gpurun_out/sanitize_${TOOL}.log
: > $OUT
TESTS=(
  "tests/test_gpu_raycast.py::test_edge_poses"
  "tests/test_gpu_raycast.py::test_fp32_sentinel_and_cap_scale_with_the_map"
  "tests/test_gpu_raycast.py::test_rotated_origin_fp64_bit_exact_and_fp32_within_tolerance[berlin--1.2]"
  "tests/test_gpu_step.py::test_opponent_raycast_random_placements"
  "tests/test_gpu_step.py::test_small_batch_split_scans_do_not_change_results[fp32-1-900]"
  "tests/test_gpu_step.py::test_small_batch_split_scans_do_not_change_results[fp64-2-120]"
  "tests/test_gpu_abi_v2.py::test_u16_scan_format_is_the_float_scan_rounded_to_half_millimetres[700]"
  "tests/test_gpu_abi_v2.py::test_u16_scan_format_is_the_float_scan_rounded_to_half_millimetres[40000]"
  "tests/test_gpu_step.py::test_rollouts_match_reference_goldens[rollout_a2_rot_m1d2]"
)
# the plain run first: the sanitizer is only started on tests that pass without it
timeout 900 python -m pytest "${TESTS[@]}" -q -x -p no:cacheprovider > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed" >> $OUT; tail -20 gpurun_out/sanitize_plain.log >> $OUT; exit 1; }
for t in "${TESTS[@]}"; do
  echo "=== $t" >> $OUT
  timeout 1200 compute-sanitizer --tool $TOOL --error-exitcode 77 --print-limit 20 \
      python -m pytest "$t" -q -x -p no:cacheprovider > gpurun_out/sanitize_one.log 2>&1
  rc=$?
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Invalid|hazard" gpurun_out/sanitize_one.log | tail -8 >> $OUT
  echo "rc=$rc" >> $OUT
  if [ $rc -ne 0 ]; then tail -60 gpurun_out/sanitize_one.log >> $OUT; fi
done
cat $OUT
