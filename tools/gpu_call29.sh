#!/bin/bash
set -u
mkdir -p gpurun_out
F110_B200_LIB=red_gym_b200/libf110_b200_phases.so timeout 600 python tools/phase_probe.py > gpurun_out/c29_phases.log 2>&1; echo "rc $?" >> gpurun_out/c29_phases.log
timeout 600 python tools/host_overhead_probe.py > gpurun_out/c29_host.log 2>&1; echo "rc $?" >> gpurun_out/c29_host.log
timeout 900 python tools/latency_probe.py > gpurun_out/c29_latency.log 2>&1; echo "rc $?" >> gpurun_out/c29_latency.log
timeout 1500 python -m pytest tests/test_gpu_step.py tests/test_gpu_abi_v2.py tests/test_gpu_dropin.py tests/test_gpu_graph.py tests/test_gpu_checked_build.py -x -q -m gpu -p no:cacheprovider > gpurun_out/c29_tests.log 2>&1; echo "rc $?" >> gpurun_out/c29_tests.log
cat gpurun_out/c29_phases.log; head -4 gpurun_out/c29_host.log; cat gpurun_out/c29_latency.log; tail -3 gpurun_out/c29_tests.log
