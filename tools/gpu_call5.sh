#!/bin/bash
# round 2, GPU call 5 (2 GPUs): the driver's multi-GPU command lines -- bench (weak + strong + NCCL gather) and the reference arm under torchrun
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 900 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/c5_bench_2gpu.json 2> gpurun_out/c5_bench_2gpu.err
echo "rc $?" >> gpurun_out/c5_bench_2gpu.err
timeout 600 $TR bench.py --impl reference --gpus 2 --steps 20 --warmup 5 > gpurun_out/c5_reference_2gpu.json 2> gpurun_out/c5_reference_2gpu.err
echo "rc $?" >> gpurun_out/c5_reference_2gpu.err
timeout 600 $TR bench.py --workload config5 --gpus 2 --steps 2 --warmup 1 --sweep-sizes 1024,4096 --sweep-beams 1080 > gpurun_out/c5_config5_2gpu.json 2> gpurun_out/c5_config5_2gpu.err
echo "rc $?" >> gpurun_out/c5_config5_2gpu.err
nvidia-smi topo -m > gpurun_out/c5_topo.txt 2>&1
ls -la gpurun_out | tail -8
