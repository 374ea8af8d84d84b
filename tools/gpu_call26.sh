#!/bin/bash
# fused kernel with 1024 threads per env (envs <= SMs): latency A/B + the fused parity tests
set -u
mkdir -p gpurun_out
timeout 900 python tools/latency_probe.py > gpurun_out/c26_latency.log 2>&1; echo "rc $?" >> gpurun_out/c26_latency.log
timeout 1200 python -m pytest tests/test_gpu_step.py tests/test_gpu_abi_v2.py tests/test_gpu_graph.py -x -q -m gpu -p no:cacheprovider > gpurun_out/c26_tests.log 2>&1; echo "rc $?" >> gpurun_out/c26_tests.log
cat gpurun_out/c26_latency.log; tail -3 gpurun_out/c26_tests.log
