#!/bin/bash
# gpurun with a retry on "no box right now" (rc 3) / transient refusals:  tools/gpurun_retry.sh <timeout s> <command...>
T=$1; shift
for i in $(seq 1 20); do
    /usr/local/graft/bin/gpurun --timeout "$T" -- "$@" > /tmp/gpurun_last.log 2>&1; rc=$?
    if [ $rc -eq 3 ] || grep -q "transient" /tmp/gpurun_last.log; then sleep 90; continue; fi
    break
done
tail -60 /tmp/gpurun_last.log; echo "gpurun rc $rc"
