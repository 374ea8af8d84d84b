#!/bin/bash
# round 2, GPU call 9 (1 GPU): records of the final tree -- whole GPU suite, smoke, default bench, launch list, one full ncu capture
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=12 -p no:cacheprovider > gpurun_out/c9_tests.log 2>&1
echo "pytest rc $?" >> gpurun_out/c9_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c9_smoke.log 2>&1
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/c9_bench.json 2> gpurun_out/c9_bench.err
echo "bench rc $?" >> gpurun_out/c9_bench.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/c9_reference.json 2> gpurun_out/c9_reference.err
timeout 300 python bench.py --workload downstream --steps 20 --warmup 5 > gpurun_out/c9_downstream.json 2>> gpurun_out/c9_bench.err
CMD="python bench.py --steps 3 --warmup 3 --preroll 60 --no-cpu --no-e2e"
$CMD > gpurun_out/c9_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_raycast32 -s 64 -c 1 -o gpurun_out/c9_raycast32 $CMD > gpurun_out/c9_ncu2.log 2>&1
ls -la gpurun_out | tail -6
