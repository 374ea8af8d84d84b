#!/bin/bash
# SM-wide round pool (fp32): bit-identity tests with the pool on, then A/B pool (F110_POOL_BELOW=100000) against the split / whole-scan launches (variant 65536)
set -u
mkdir -p gpurun_out; rm -f gpurun_out/c20_pool.jsonl
timeout 900 python -m pytest tests/test_gpu_step.py tests/test_gpu_abi_v2.py tests/test_gpu_edge_cases.py tests/test_gpu_graph.py -m gpu -q -x -p no:cacheprovider -k "small_batch or step_host or u16 or beam_counts or graph or masked or auto_reset or fp32" > gpurun_out/c20_tests.log 2>&1; echo "rc $?" >> gpurun_out/c20_tests.log
tail -5 gpurun_out/c20_tests.log
F110_POOL_BELOW=100000 timeout 600 python -m pytest tests/test_gpu_step.py -m gpu -q -x -p no:cacheprovider -k "small_batch or step_host_matches or sharded" > gpurun_out/c20_tests_big.log 2>&1; echo "rc $?" >> gpurun_out/c20_tests_big.log
tail -3 gpurun_out/c20_tests_big.log
for n in 600 1500 3000 4096 6144 8192 12288; do for v in 65536 0 65536 0; do
  F110_POOL_BELOW=100000 timeout 300 python bench.py --workload config2 --dtype fp32 --envs $n --steps 100 --warmup 10 --no-cpu --no-e2e --no-graph --variant $v >> gpurun_out/c20_pool.jsonl 2>> gpurun_out/c20_pool.err
done; done
python - <<'PY'
import json
for l in open('gpurun_out/c20_pool.jsonl'):
    if l.startswith('{'):
        d=json.loads(l); print(d['config']['envs_per_gpu'], 'pool' if d['config']['raycast_variant']==0 else 'split/whole', 'ms/step %.4f raycast %.4f'%(d['ms_per_step'], d['roofline']['launch_ms']))
PY
