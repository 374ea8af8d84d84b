#!/usr/bin/env python
"""Step rate of small batches, eager f110_step calls against CUDA-graph replays (BatchF110Env.capture_step_graph, 16 steps per
graph), actions resident on the device.   python tools/graph_probe.py"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from red_gym_b200 import BatchF110Env, MAPS_DIR
from red_gym_b200.host import raceline_poses

K = 16
for dtype, n in (("fp64", 1), ("fp64", 64), ("fp64", 296), ("fp64", 1024), ("fp64", 4096), ("fp32", 1024), ("fp32", 4096), ("fp32", 16384)):
    env = BatchF110Env(num_envs=n, num_agents=1, map=os.path.join(MAPS_DIR, "example_map"), dtype=dtype, auto_reset=True,
                       reset_poses=raceline_poses(256)[:, None, :], noise="philox" if dtype == "fp32" else "numpy")
    env.reset(raceline_poses(n)[:, None, :])
    g = torch.Generator(device="cuda").manual_seed(1)
    acts = torch.rand((K, n, 1, 2), generator=g, device="cuda", dtype=torch.float64)
    acts[..., 0] = (acts[..., 0] * 2 - 1) * 0.4189
    acts[..., 1] *= 8.0
    for i in range(200):
        env.step(acts[i % K])
    torch.cuda.synchronize()
    R = 60
    t0 = time.perf_counter()
    for r in range(R):
        for k in range(K):
            env.step(acts[k])
    torch.cuda.synchronize()
    eager = (time.perf_counter() - t0) / (R * K) * 1e6
    env.reserve_noise(K * (R + 8) + 8)
    graph = env.capture_step_graph(acts, steps=K, warmup=3)
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for r in range(R):
        graph.replay()
    torch.cuda.synchronize()
    rep = (time.perf_counter() - t0) / (R * K) * 1e6
    print("%s %6d envs: eager %.1f us per step (%.3g agent-steps/s), graph replay %.1f us per step (%.3g agent-steps/s), x %.2f"
          % (dtype, n, eager, n / eager * 1e6, rep, n / rep * 1e6, eager / rep))
    env.close()
