#!/bin/bash
# round 2, GPU call 6 (1 GPU): smoke(), A/B probes (round stealing, scans per CTA), full suite on the current tree
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c6_smoke.log 2>&1
echo "smoke rc $?" >> gpurun_out/c6_smoke.log
for v in 0 1024 8192 0 1024 8192; do
  timeout 300 python bench.py --steps 10 --warmup 3 --preroll 150 --no-cpu --no-e2e --variant $v >> gpurun_out/c6_wpc.jsonl 2>> gpurun_out/c6_wpc.err
done
python tools/steal_probe.py > gpurun_out/c6_steal.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q --maxfail=12 -p no:cacheprovider > gpurun_out/c6_tests.log 2>&1
echo "pytest rc $?" >> gpurun_out/c6_tests.log
ls -la gpurun_out | tail -6
