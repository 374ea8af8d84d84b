#!/bin/bash
set -u
mkdir -p gpurun_out
F110_B200_LIB=red_gym_b200/libf110_b200_phases.so timeout 600 python tools/phase_probe.py > gpurun_out/c27_phases.log 2>&1; echo "rc $?" >> gpurun_out/c27_phases.log
cat gpurun_out/c27_phases.log
