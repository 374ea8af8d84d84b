#!/bin/bash
# round 2, GPU call 2: the tests that failed / are new, default bench, launch list + one full ncu capture of the bench-size ray-cast
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_abi_v2.py tests/test_gpu_raycast.py tests/test_gpu_step.py tests/test_gpu_dropin.py -m gpu -q --maxfail=12 -s -p no:cacheprovider > gpurun_out/c2_tests.log 2>&1
echo "pytest rc $?" >> gpurun_out/c2_tests.log
python tools/latency_probe.py > gpurun_out/c2_latency.log 2>&1
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/c2_bench.json 2> gpurun_out/c2_bench.err
echo "bench rc $?" >> gpurun_out/c2_bench.err
timeout 300 python bench.py --workload config2 --steps 200 --warmup 20 --no-cpu > gpurun_out/c2_bench_config2.json 2>> gpurun_out/c2_bench.err
timeout 300 python bench.py --workload config4 --steps 50 --warmup 10 --no-cpu > gpurun_out/c2_bench_config4.json 2>> gpurun_out/c2_bench.err
CMD="python bench.py --steps 3 --warmup 3 --preroll 60 --no-cpu --no-e2e"
$CMD > gpurun_out/c2_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 60 --csv --log-file gpurun_out/c2_launches.csv $CMD > gpurun_out/c2_ncu1.log 2>&1
$CMD > gpurun_out/c2_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_raycast32 -s 64 -c 1 -o gpurun_out/c2_raycast32 $CMD > gpurun_out/c2_ncu2.log 2>&1
ls -la gpurun_out
