#!/bin/bash
# round 2, GPU call 4: checked build run, latency after the noise fix, mid-size lines, config 5 sweep + its ncu metrics pass (1 GPU)
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_checked_build.py tests/test_gpu_dropin.py tests/test_gpu_abi_v2.py -m gpu -q -s -p no:cacheprovider > gpurun_out/c4_tests.log 2>&1
echo "pytest rc $?" >> gpurun_out/c4_tests.log
python tools/latency_probe.py > gpurun_out/c4_latency.log 2>&1
timeout 300 python examples/waypoint_follow.py > gpurun_out/c4_waypoint.log 2>&1
timeout 300 python bench.py --workload config2 --steps 200 --warmup 20 --no-cpu > gpurun_out/c4_bench_config2.json 2> gpurun_out/c4_bench.err
timeout 300 python bench.py --workload config4 --steps 50 --warmup 10 --no-cpu > gpurun_out/c4_bench_config4.json 2>> gpurun_out/c4_bench.err
CMD="python bench.py --workload config5 --steps 2 --warmup 1"
timeout 280 $CMD > gpurun_out/c4_config5_sweep_1gpu.json 2>> gpurun_out/c4_bench.err &&
ncu --metrics gpu__time_duration.sum,lts__t_sector_hit_rate.pct,lts__t_sectors.sum,l1tex__m_xbar2l1tex_read_bytes.sum,dram__bytes_read.sum,dram__bytes_write.sum,l1tex__t_sector_hit_rate.pct \
    --clock-control none -k regex:'k_scan_only64|k_raycast32' --csv --log-file gpurun_out/c4_config5_ncu.csv $CMD > gpurun_out/c4_ncu.log 2>&1
ls -la gpurun_out
