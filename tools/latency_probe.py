#!/usr/bin/env python
"""Single-env step latency breakdown: wall time of f110_step_host, GPU time of the whole step (CUDA events around the
call), GPU time of the ray-cast launch alone.   python tools/latency_probe.py"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from red_gym_b200 import BatchF110Env, MAPS_DIR

# variant 256: the three-kernel step (k_dynamics / k_raycast64 / k_finalize + k_export_host) instead of k_step_fused64
# variant 1: the fused kernel with 256 threads per env even when there are fewer envs than SMs (default: 1024 then)
for n, yaw, variant in ((1, 1.37079632679, 0), (1, 1.37079632679, 1), (1, 1.37079632679, 256), (64, 1.37079632679, 1), (148, 1.37079632679, 0), (148, 1.37079632679, 1), (1, 0.0, 0), (64, 1.37079632679, 0), (64, 1.37079632679, 256), (296, 1.37079632679, 0),
                        (296, 1.37079632679, 256), (4096, 1.37079632679, 0), (4096, 0.0, 0)):
    # yaw == 0 exactly is the yaw of every car stopped by iTTC: the last beam of its scans replays the reference's running index sum
    env = BatchF110Env(num_envs=n, num_agents=1, map=os.path.join(MAPS_DIR, "example_map"), dtype="fp64", noise="numpy", raycast_variant=variant)
    env.reset(np.tile(np.array([[[0.7, 0.0, yaw]]]), (n, 1, 1)))
    act = torch.zeros((n, 1, 2), dtype=torch.float64).pin_memory(); act[..., 1] = 0.0
    obs = torch.zeros((10, n, 1), dtype=torch.float64).pin_memory()
    scans = torch.zeros((n, 1, 1080), dtype=torch.float64).pin_memory()
    done = torch.zeros((n,), dtype=torch.uint8).pin_memory()
    actd = act.cuda()
    env.reserve_noise(8192)       # parity-mode noise rows for the whole probe: their generation (numpy) is not part of the call timed here
    for _ in range(200):
        env.step_host(act, obs, scans, done)
    K = 2000
    t0 = time.perf_counter()
    for _ in range(K):
        env.step_host(act, obs, scans, done)
    wall = (time.perf_counter() - t0) / K * 1e6
    env.enable_timing(True)       # from here on the library records two events per step (last_raycast_ms)
    rc = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    gpu = []
    for _ in range(200):
        e0.record(); env.step(actd); e1.record(); torch.cuda.synchronize()
        gpu.append(e0.elapsed_time(e1) * 1e3); rc.append(env.last_raycast_ms() * 1e3)
    t0 = time.perf_counter()
    for _ in range(K):
        env.step(actd)
    torch.cuda.synchronize()
    dev = (time.perf_counter() - t0) / K * 1e6
    print("envs %d yaw %g variant %d: step_host wall %.1f us | device step back-to-back %.1f us per step | one step GPU (events) %.1f us, ray-cast launch %.1f us"
          % (n, yaw, variant, wall, dev, float(np.median(gpu)), float(np.median(rc))))
