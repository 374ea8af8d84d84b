#!/bin/bash
# the whole GPU suite against the bounds-checked build of the final tree
set -u
mkdir -p gpurun_out
timeout 2400 python tools/checked_run.py tests > gpurun_out/c36_checked_suite.log 2>&1; echo "rc $?" >> gpurun_out/c36_checked_suite.log
grep "CHECKED RUN" gpurun_out/c36_checked_suite.log; tail -4 gpurun_out/c36_checked_suite.log
