#!/usr/bin/env python
"""Ray-cast throughput per bundled map (free-space poses, scan-only kernels): the fp64 path uses a one-instruction cell
index on maps whose resolution is a power of two (example_map) and a longer one elsewhere (0.05 m maps).
python tools/map_probe.py [n]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from red_gym_b200 import BatchF110Env, MAPS_DIR

n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
for name in ("example_map", "vegas", "berlin", "skirk", "stata_basement"):
    for dtype in ("fp64", "fp32"):
        env = BatchF110Env(num_envs=1, num_agents=1, map=os.path.join(MAPS_DIR, name), dtype=dtype, noise="off", edt="scipy")
        m = env.map
        free = np.argwhere(m.edt > 0.3)
        rng = np.random.default_rng(7)
        pick = free[rng.choice(free.shape[0], n)]
        x = m.origin[0] + (pick[:, 1] + rng.uniform(0, 1, n)) * m.resolution
        y = m.origin[1] + (pick[:, 0] + rng.uniform(0, 1, n)) * m.resolution
        p = torch.as_tensor(np.stack([x, y, rng.uniform(0, 2 * np.pi, n)], axis=1), device="cuda")
        out = env.raycast(p)
        env.count_lookups(True); env.raycast(p, out=out); lk = env.lookups() / n; env.count_lookups(False)
        for _ in range(3):
            env.raycast(p, out=out)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            env.raycast(p, out=out)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print("%-15s %s res %.4f  %d x %d  lookups/scan %.0f  %.3f ms  %.2f ns/scan  %.2f ps/lookup" %
              (name, dtype, m.resolution, m.height, m.width, lk, ms, 1e6 * ms / n, 1e9 * ms / n / lk))
        del env
