#!/bin/bash
# round 2, GPU call 8 (1 GPU): scan records prepared by the sort -- A/B against in-kernel preparation (variant 16384), full suite
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=12 -p no:cacheprovider > gpurun_out/c8_tests.log 2>&1
echo "pytest rc $?" >> gpurun_out/c8_tests.log
rm -f gpurun_out/c8_rec.jsonl
for v in 0 16384 0 16384; do
  timeout 300 python bench.py --steps 20 --warmup 3 --preroll 150 --no-cpu --no-e2e --variant $v >> gpurun_out/c8_rec.jsonl 2>> gpurun_out/c8_rec.err
done
for n in 65536 262144; do for v in 0 16384; do
  timeout 300 python bench.py --steps 30 --warmup 3 --preroll 150 --no-cpu --no-e2e --envs $n --variant $v >> gpurun_out/c8_rec.jsonl 2>> gpurun_out/c8_rec.err
done; done
CMD="python bench.py --steps 3 --warmup 3 --preroll 60 --no-cpu --no-e2e"
$CMD > gpurun_out/c8_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 60 --csv --log-file gpurun_out/c8_launches.csv $CMD > gpurun_out/c8_ncu1.log 2>&1
ls -la gpurun_out | tail -5
