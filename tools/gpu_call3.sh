#!/bin/bash
# round 2, GPU call 3: whole GPU suite on the current tree, latency probe, mid-size bench lines, then compute-sanitizer memcheck
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --maxfail=12 -p no:cacheprovider > gpurun_out/c3_tests.log 2>&1
echo "pytest rc $?" >> gpurun_out/c3_tests.log
python tools/latency_probe.py > gpurun_out/c3_latency.log 2>&1
timeout 300 python bench.py --workload config2 --steps 200 --warmup 20 --no-cpu > gpurun_out/c3_bench_config2.json 2> gpurun_out/c3_bench.err
timeout 300 python bench.py --workload config2 --steps 200 --warmup 20 --no-cpu --no-e2e --dtype fp32 > gpurun_out/c3_bench_config2_fp32.json 2>> gpurun_out/c3_bench.err
timeout 300 python examples/waypoint_follow.py > gpurun_out/c3_waypoint.log 2>&1
bash tools/sanitize.sh memcheck > gpurun_out/c3_sanitize_stdout.log 2>&1
ls -la gpurun_out
