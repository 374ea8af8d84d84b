#!/bin/bash
# A/B: spatial sort forced on (variant 32768) below F110_ORDER_FROM, steady state (200 untimed steps), fp64 and fp32
set -u
mkdir -p gpurun_out; rm -f gpurun_out/c12_sort.jsonl
for dt in fp64 fp32; do for n in 4096 8192 16384 24576; do for v in 0 32768 0 32768; do
  timeout 300 python bench.py --workload config2 --dtype $dt --envs $n --steps 100 --warmup 10 --no-cpu --no-e2e --variant $v >> gpurun_out/c12_sort.jsonl 2>> gpurun_out/c12_sort.err
done; done; done
python - <<'PY'
import json
for l in open('gpurun_out/c12_sort.jsonl'):
    if l.startswith('{'):
        d=json.loads(l); print(d['dtype'], d['config']['envs_per_gpu'], 'variant', d['config']['raycast_variant'], 'ms/step %.4f raycast %.4f'%(d['ms_per_step'], d['roofline']['launch_ms']))
PY
