#!/bin/bash
# round 2, GPU call 1: GPU test suite, default bench, launch list + one full ncu capture of the bench-size ray-cast
set -u
mkdir -p gpurun_out
python -c "import torch; print(torch.cuda.get_device_name(0))" > gpurun_out/c1_env.log 2>&1
nproc >> gpurun_out/c1_env.log; grep -m1 "model name" /proc/cpuinfo >> gpurun_out/c1_env.log
timeout 1500 python -m pytest tests -m gpu -q --maxfail=12 -s -p no:cacheprovider > gpurun_out/c1_tests.log 2>&1
echo "pytest rc $?" >> gpurun_out/c1_tests.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err
echo "bench rc $?" >> gpurun_out/c1_bench.err
CMD="python bench.py --steps 3 --warmup 3 --preroll 60 --no-cpu --no-e2e"
$CMD > gpurun_out/c1_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 60 --csv --log-file gpurun_out/c1_launches.csv $CMD > gpurun_out/c1_ncu1.log 2>&1
$CMD > gpurun_out/c1_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_raycast32 -s 64 -c 1 -o gpurun_out/c1_raycast32 $CMD > gpurun_out/c1_ncu2.log 2>&1
ls -la gpurun_out
