#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/c30_suite.log 2>&1; echo "suite rc $?" >> gpurun_out/c30_suite.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c30_smoke.log 2>&1; echo "smoke rc $?" >> gpurun_out/c30_smoke.log
tail -4 gpurun_out/c30_suite.log; tail -2 gpurun_out/c30_smoke.log
