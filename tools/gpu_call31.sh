#!/bin/bash
# launch list of the final tree's default bench (ncu gpu__time_duration pass), after the same command ran clean without ncu
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 5 --warmup 3 --preroll 60 --no-cpu --no-e2e"
$CMD > gpurun_out/c31_plain.json 2> gpurun_out/c31_plain.err; echo "plain rc $?" >> gpurun_out/c31_plain.err
ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 60 --csv --log-file gpurun_out/c31_launches.csv $CMD > gpurun_out/c31_ncu.log 2>&1; echo "ncu rc $?" >> gpurun_out/c31_ncu.log
python tools/launch_summary.py gpurun_out/c31_launches.csv > gpurun_out/c31_launch_summary.txt 2>&1
cat gpurun_out/c31_launch_summary.txt; tail -1 gpurun_out/c31_plain.err; tail -1 gpurun_out/c31_ncu.log
