#!/bin/bash
# no-alloc entry point: its test alone in a fresh process, then the whole GPU suite (-x, as the round's end runs it) with it last
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -p no:cacheprovider > gpurun_out/c10_alone.log 2>&1
echo "alone rc $?" >> gpurun_out/c10_alone.log
timeout 1800 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/c10_suite.log 2>&1
echo "suite rc $?" >> gpurun_out/c10_suite.log
tail -4 gpurun_out/c10_alone.log; tail -4 gpurun_out/c10_suite.log
