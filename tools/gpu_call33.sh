#!/bin/bash
# 8 GPUs, final tree: the bench line at N = 8 (weak + strong + NCCL gather + e2e)
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
timeout 900 $TR bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu > gpurun_out/c33_bench_8gpu.json 2> gpurun_out/c33_bench_8gpu.err
echo "rc $?" >> gpurun_out/c33_bench_8gpu.err
tail -3 gpurun_out/c33_bench_8gpu.err; tail -c 1500 gpurun_out/c33_bench_8gpu.json
