#!/bin/bash
# final verification of the tree: GPU suite (-x, as the round's end runs it), smoke, default bench, reference arm
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/c19_suite.log 2>&1; echo "suite rc $?" >> gpurun_out/c19_suite.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c19_smoke.log 2>&1; echo "smoke rc $?" >> gpurun_out/c19_smoke.log
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/c19_reference.json 2> gpurun_out/c19_reference.err; echo "ref rc $?" >> gpurun_out/c19_reference.err
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/c19_bench.json 2> gpurun_out/c19_bench.err; echo "bench rc $?" >> gpurun_out/c19_bench.err
tail -4 gpurun_out/c19_suite.log; tail -2 gpurun_out/c19_smoke.log; tail -1 gpurun_out/c19_reference.err; tail -1 gpurun_out/c19_bench.err
