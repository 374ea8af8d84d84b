#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python tools/checked_run.py tests/test_gpu_edge_cases.py > gpurun_out/c38_edge_checked.log 2>&1; echo "rc $?" >> gpurun_out/c38_edge_checked.log
grep "CHECKED RUN" gpurun_out/c38_edge_checked.log | tail -1; tail -3 gpurun_out/c38_edge_checked.log
timeout 2400 python tools/checked_run.py tests > gpurun_out/c38_checked_suite.log 2>&1; echo "rc $?" >> gpurun_out/c38_checked_suite.log
grep "CHECKED RUN" gpurun_out/c38_checked_suite.log | tail -1; tail -3 gpurun_out/c38_checked_suite.log
timeout 2400 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/c38_suite.log 2>&1; echo "suite rc $?" >> gpurun_out/c38_suite.log
tail -3 gpurun_out/c38_suite.log
