#!/usr/bin/env python
"""Runs GPU tests against the bounds-checked build of the library (libf110_b200_checked.so, -DF110_CHECKED) and fails if
any of its checks fired.  compute-sanitizer is closed on the B200 pool, so the kernels carry their own tests of every
address / index whose bounds test the fast path removed or hoisted (csrc/f110_dev.h: F110_CHECK).
    python tools/checked_run.py [pytest node ids ...]          (default: the list of tools/sanitize.sh)
Prints one line `CHECKED RUN: ...` with the counters; exit code 0 only if the tests passed, the self-test violation was
counted (the checks are compiled in) and every other counter is zero."""
import ctypes
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(REPO, "red_gym_b200", "libf110_b200_checked.so")
DEFAULT = [
    "tests/test_gpu_raycast.py::test_edge_poses",
    "tests/test_gpu_raycast.py::test_fp32_sentinel_and_cap_scale_with_the_map",
    "tests/test_gpu_raycast.py::test_rotated_origin_fp64_bit_exact_and_fp32_within_tolerance",
    "tests/test_gpu_raycast.py::test_fp64_bit_exact_vs_oracle_random_poses",
    "tests/test_gpu_raycast.py::test_raycast_spatial_order_is_only_a_schedule",
    "tests/test_gpu_step.py::test_opponent_raycast_random_placements",
    "tests/test_gpu_step.py::test_small_batch_split_scans_do_not_change_results",
    "tests/test_gpu_step.py::test_spatial_launch_order_does_not_change_results",
    "tests/test_gpu_step.py::test_fused_step_is_bit_identical_to_the_three_kernel_step",
    "tests/test_gpu_step.py::test_rollouts_match_reference_goldens",
    "tests/test_gpu_step.py::test_fp32_two_agent_rollout_tracks_fp64_path",
    "tests/test_gpu_abi_v2.py::test_u16_scan_format_is_the_float_scan_rounded_to_half_millimetres",
    "tests/test_gpu_edge_cases.py::test_raycast_beam_count_edges",
    "tests/test_gpu_edge_cases.py::test_step_with_other_beam_counts",
]
NAMES = ["edt32_address", "edt64_cell", "table_index", "queue_slot", "order_slot", "beam_index", "noise_row", "selftest"]


def main():
    os.environ["F110_B200_LIB"] = LIB
    sys.path.insert(0, REPO)
    import pytest
    from red_gym_b200 import _native as nat
    L = nat.load()
    assert L.f110_build_is_checked() == 1, "not the checked build: %s" % nat.LIB_PATH
    tests = sys.argv[1:] or DEFAULT
    rc = pytest.main(["-q", "-x", "-p", "no:cacheprovider", "-m", "gpu"] + [os.path.join(REPO, t) for t in tests])
    out = (ctypes.c_uint64 * 8)()
    nat.check(L.f110_debug_violations(0, out, 1))
    counts = dict(zip(NAMES, [int(v) for v in out]))
    ok = int(rc) == 0 and counts["selftest"] == 1 and all(v == 0 for k, v in counts.items() if k != "selftest")
    print("CHECKED RUN: pytest rc %d, violations %s -> %s" % (int(rc), counts, "clean" if ok else "FAILED"))
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
