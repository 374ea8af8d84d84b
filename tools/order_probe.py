#!/usr/bin/env python
"""Diagnostic: how fine must the spatial sort key be?  Steps a config-3 env until env order has decorrelated from
track order, takes its poses, and times f110_raycast on them (device reorder off, variant 64) in host-sorted orders
of different key resolutions.   python tools/order_probe.py [n ...]"""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch
from red_gym_b200 import BatchF110Env, MAPS_DIR
from red_gym_b200.host import raceline_poses

m = os.path.join(MAPS_DIR, "example_map")
sizes = [int(a) for a in sys.argv[1:]] or [32768, 262144]
for n in sizes:
    env = BatchF110Env(num_envs=n, num_agents=1, map=m, dtype="fp32", noise="philox", auto_reset=True,
                       reset_poses=raceline_poses(1024)[:, None, :], raycast_variant=64)
    env.reset(raceline_poses(n)[:, None, :])
    g = torch.Generator(device="cuda").manual_seed(1)
    for t in range(300):
        a = torch.rand((n, 1, 2), generator=g, device="cuda", dtype=torch.float64)
        a[..., 0] = (a[..., 0] * 2 - 1) * 0.4189; a[..., 1] *= 8.0
        env.step(a)
    st = env.state.cpu().numpy()          # [7][n][1]
    p = np.stack([st[0, :, 0], st[1, :, 0], st[4, :, 0]], axis=1)
    res = env.map.resolution; ox, oy = env.map.origin[0], env.map.origin[1]
    gx, gy = (p[:, 0] - ox) / res, (p[:, 1] - oy) / res
    def keyed(tile_cells, nh, morton=False):
        tx, ty = np.floor(gx / tile_cells).astype(np.int64), np.floor(gy / tile_cells).astype(np.int64)
        hb = np.floor(p[:, 2] / (2 * np.pi) * nh).astype(np.int64) % nh
        return np.lexsort((hb, tx, ty))
    orders = {"as-is": np.arange(n), "2m/8": keyed(32, 8), "2m/16": keyed(32, 16), "1m/8": keyed(16, 8), "1m/16": keyed(16, 16),
              "0.5m/16": keyed(8, 16), "0.5m/32": keyed(8, 32), "0.25m/32": keyed(4, 32), "0.25m/64": keyed(4, 64)}
    out = None
    line = []
    for name, o in orders.items():
        pp = torch.as_tensor(p[o], device="cuda")
        out = env.raycast(pp, out=out)
        for _ in range(2):
            env.raycast(pp, out=out)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            env.raycast(pp, out=out)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        line.append("%s %.4f ms (%.2f ns)" % (name, ms, 1e6 * ms / n))
    print(n, " | ".join(line))
    del env
