#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 300 python bench.py --workload config2 --steps 200 --warmup 20 --no-cpu > gpurun_out/c17_config2.json 2> gpurun_out/c17.err; echo "rc $?" >> gpurun_out/c17.err
timeout 300 python bench.py --workload config4 --steps 48 --warmup 10 --no-cpu > gpurun_out/c17_config4.json 2>> gpurun_out/c17.err; echo "rc $?" >> gpurun_out/c17.err
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/c17_default.json 2>> gpurun_out/c17.err; echo "rc $?" >> gpurun_out/c17.err
grep "^rc\|Error\|error" gpurun_out/c17.err | tail -5
python - <<'PY'
import json
for f in ('c17_config2','c17_config4','c17_default'):
    d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
    print(f, 'value %.4g ms/step %.4f'%(d['value'], d['ms_per_step']), 'graph', d.get('graph_replay') and {k:(round(v,4) if isinstance(v,float) else v) for k,v in d['graph_replay'].items() if k!='api'})
PY
