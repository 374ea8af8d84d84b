#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 2400 python tools/checked_run.py tests > gpurun_out/c39_checked_suite.log 2>&1; echo "rc $?" >> gpurun_out/c39_checked_suite.log
grep "CHECKED RUN" gpurun_out/c39_checked_suite.log | tail -1; tail -3 gpurun_out/c39_checked_suite.log
timeout 1200 python -m pytest tests/test_gpu_bitmap.py tests/test_gpu_pursuit.py tests/test_gpu_render.py tests/test_gpu_checked_build.py -x -q -m gpu -p no:cacheprovider > gpurun_out/c39_tests.log 2>&1; echo "rc $?" >> gpurun_out/c39_tests.log
tail -3 gpurun_out/c39_tests.log
timeout 300 python bench.py --workload downstream --steps 20 --warmup 5 > gpurun_out/c39_downstream.json 2> gpurun_out/c39_downstream.err; echo "rc $?" >> gpurun_out/c39_downstream.err
tail -c 900 gpurun_out/c39_downstream.json
