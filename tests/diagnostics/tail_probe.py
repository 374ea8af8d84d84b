#!/usr/bin/env python
"""Diagnostic (uses the oracle to count lookups per scan, so it is NOT product code): how much of the mid-size batch
inefficiency of the ray-cast is the spread of per-scan cost?  Times f110_raycast over the same set of raceline poses
(with lateral/heading jitter) in natural, random and longest-first order.   python tests/diagnostics/tail_probe.py"""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "oracle"))
import torch
import f110_oracle as fo
from red_gym_b200 import BatchF110Env, MAPS_DIR
from red_gym_b200.host import raceline_poses

m = os.path.join(MAPS_DIR, "example_map")
for dtype in ("fp32", "fp64"):
    env = BatchF110Env(num_envs=1, num_agents=1, map=m, dtype=dtype, noise="off", raycast_variant=64)   # 64: no spatial reorder
    ora = fo.OracleEnv(1, 1, map=m, noise=False, nthreads=os.cpu_count())
    for n in (8192, 16384, 32768, 65536):
        rng = np.random.default_rng(1)
        p = raceline_poses(n)
        p[:, 2] += rng.uniform(-0.3, 0.3, n)
        _, lk = ora.scan(p, return_lookups=True)
        lk = lk.astype(np.int64)
        orders = {"natural": np.arange(n), "random": rng.permutation(n), "longest-first": np.argsort(-lk, kind="stable"),
                  "shortest-first": np.argsort(lk, kind="stable")}
        res = {}
        for name, o in orders.items():
            pp = torch.as_tensor(p[o], device="cuda")
            out = env.raycast(pp)
            for _ in range(3):
                env.raycast(pp, out=out)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                env.raycast(pp, out=out)
            e1.record(); torch.cuda.synchronize()
            res[name] = e0.elapsed_time(e1) / 20
        print(dtype, n, "lookups/scan mean %.0f min %d max %d" % (lk.mean(), lk.min(), lk.max()),
              "  ".join("%s %.4f ms (%.2f ns/scan)" % (k, v, 1e6 * v / n) for k, v in res.items()))
