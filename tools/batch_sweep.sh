#!/bin/bash
# Batch-size sweep of the step (config 3 workload) over raycast variants (default: 0), both dtypes.
# usage: VARIANTS="0 128" tools/batch_sweep.sh out.jsonl [sizes...]
out=$1; shift
sizes=${@:-"4096 8192 16384 32768 65536 131072"}
: > $out
for dt in fp32 fp64; do for n in $sizes; do for v in ${VARIANTS:-0}; do
  line=$(timeout 120 python bench.py --envs $n --dtype $dt --variant $v --no-cpu --no-e2e --steps 40 --warmup 10 2>/dev/null | grep '^{' | tail -1)
  echo "{\"dtype\": \"$dt\", \"envs\": $n, \"variant\": $v, \"line\": $line}" >> $out
  python - "$line" $dt $n $v <<'PY'
import json, sys
j = json.loads(sys.argv[1])
print(sys.argv[2], sys.argv[3], "variant", sys.argv[4], "ms/step %.4f" % j["ms_per_step"], "raycast ms %.4f" % j["roofline"]["launch_ms"], "ns/scan %.2f" % (1e6 * j["ms_per_step"] / int(sys.argv[3])))
PY
done; done; done
