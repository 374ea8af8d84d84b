#!/usr/bin/env python
"""A/B of round stealing inside split CTAs (raycast_variant bit 9 = off) on one box: ray-cast launch time of the steady-state
config-2-like workload at several batch sizes, fp64 and fp32.   python tools/steal_probe.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from red_gym_b200 import BatchF110Env, MAPS_DIR
from red_gym_b200.host import raceline_poses

for dtype in ("fp64", "fp32"):
    for n in (600, 1500, 3000, 4096, 8192, 16384):
        if dtype == "fp32" and n > 4500:
            continue
        res = []
        for variant in (0, 512, 0, 512):
            env = BatchF110Env(num_envs=n, num_agents=1, map=os.path.join(MAPS_DIR, "example_map"), dtype=dtype, auto_reset=True,
                               reset_poses=raceline_poses(1024)[:, None, :], raycast_variant=variant)
            env.reset(raceline_poses(n)[:, None, :])
            g = torch.Generator(device="cuda").manual_seed(5)
            acts = []
            for _ in range(8):
                a = torch.rand((n, 1, 2), generator=g, device="cuda", dtype=torch.float64)
                a[..., 0] = (a[..., 0] * 2 - 1) * 0.4189
                a[..., 1] *= 8.0
                acts.append(a)
            for i in range(220):
                env.step(acts[i % 8])
            env.enable_timing(True)
            ms = []
            for i in range(100):
                env.step(acts[i % 8])
                ms.append(env.last_raycast_ms())
            res.append(float(np.median(ms)) * 1e3)
            env.close()
        print("%s %6d envs: ray-cast launch, stealing on %.1f / %.1f us, off %.1f / %.1f us" % (dtype, n, res[0], res[2], res[1], res[3]))
