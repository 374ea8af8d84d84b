#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_abi_v2.py tests/test_gpu_step.py tests/test_gpu_graph.py tests/test_gpu_dropin.py -x -q -m gpu -p no:cacheprovider > gpurun_out/c32_tests.log 2>&1; echo "rc $?" >> gpurun_out/c32_tests.log
for i in 1 2; do timeout 300 python bench.py --workload config2 --steps 400 --warmup 20 --no-cpu > gpurun_out/c32_config2_$i.json 2> gpurun_out/c32_$i.err; echo "rc $?" >> gpurun_out/c32_$i.err; done
tail -3 gpurun_out/c32_tests.log
python - <<'PY'
import json
for i in (1, 2):
    d=json.loads(open('gpurun_out/c32_config2_%d.json'%i).read().strip().splitlines()[-1])
    print('config2 run', i, 'value %.4g ms/step %.4f'%(d['value'], d['ms_per_step']), 'raycast ms', d['roofline'].get('launch_ms'), 'graph', d.get('graph_replay') and d['graph_replay']['ms_per_step'])
PY
