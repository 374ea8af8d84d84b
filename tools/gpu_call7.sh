#!/bin/bash
# round 2, GPU call 7 (8 GPUs): BASELINE config 5 on 8 x B200 and the bench line at N = 8 (weak + strong + NCCL gather + e2e roof)
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519"
timeout 700 $TR bench.py --workload config5 --gpus 8 --steps 3 --warmup 3 > gpurun_out/c7_config5_8gpu.json 2> gpurun_out/c7_config5_8gpu.err
echo "rc $?" >> gpurun_out/c7_config5_8gpu.err
timeout 900 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/c7_bench_8gpu.json 2> gpurun_out/c7_bench_8gpu.err
echo "rc $?" >> gpurun_out/c7_bench_8gpu.err
nvidia-smi topo -m > gpurun_out/c7_topo.txt 2>&1
nproc >> gpurun_out/c7_topo.txt
ls -la gpurun_out | tail -5
