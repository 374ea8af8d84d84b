#!/bin/bash
# verification of the tree: GPU suite (-x, as the round's end runs it), smoke, reference arm, default bench
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/c54_suite.log 2>&1; echo "suite rc $?" >> gpurun_out/c54_suite.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c54_smoke.log 2>&1; echo "smoke rc $?" >> gpurun_out/c54_smoke.log
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/c54_reference.json 2> gpurun_out/c54_reference.err; echo "ref rc $?" >> gpurun_out/c54_reference.err
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/c54_bench.json 2> gpurun_out/c54_bench.err; echo "bench rc $?" >> gpurun_out/c54_bench.err
tail -3 gpurun_out/c54_suite.log; tail -2 gpurun_out/c54_smoke.log; tail -1 gpurun_out/c54_reference.err; tail -1 gpurun_out/c54_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/c54_bench.json').read().strip().splitlines()[-1])
r=json.loads(open('gpurun_out/c54_reference.json').read().strip().splitlines()[-1])
print('value %.4g ms/step %.3f e2e %.4g frac %.3f compact %.4g cpu_baseline %.4g %s launches %s clocks %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e'].get('frac_of_pcie_peak', 0), d['e2e']['compact_u16']['value'], d['cpu_baseline']['value'], d['cpu_baseline']['kind'], d['gpu_launches'], d['clocks']))
print('reference arm %.4g %s cores %s' % (r['value'], r['cpu_baseline']['kind'], r['cpu_baseline']['cores']))
PY
