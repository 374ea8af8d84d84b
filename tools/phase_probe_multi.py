#!/usr/bin/env python
"""Phases of the fused multi-agent single-env step (k_step_fused_multi64), clock64 stamps of thread 0 (build.py --phases).
   F110_B200_LIB=red_gym_b200/libf110_b200_phases.so python tools/phase_probe_multi.py"""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from red_gym_b200 import BatchF110Env, MAPS_DIR, _native
lib = _native.load()
lib.f110_debug_phases.argtypes = [ctypes.c_void_p]
names = ["dynamics + prep + rounds of warp 0", "barrier: other warps' rounds", "iTTC + pair records + GJK + barrier", "walk all records (+ full batches)",
         "last batch cast", "barrier", "lap bookkeeping + export"]
order = [0, 3, 4, 1, 2, 5, 7, 6]
for A in (2, 4):
    env = BatchF110Env(num_envs=1, num_agents=A, map=os.path.join(MAPS_DIR, "example_map"), dtype="fp64", noise="numpy")
    env.reset(np.array([[[0.7 + 1.2 * a, 0.0, 1.37079632679] for a in range(A)]]))
    act = torch.zeros((1, A, 2), dtype=torch.float64).pin_memory(); act[..., 1] = 1.0
    obs = torch.zeros((10, 1, A), dtype=torch.float64).pin_memory()
    scans = torch.zeros((1, A, 1080), dtype=torch.float64).pin_memory()
    done = torch.zeros((1,), dtype=torch.uint8).pin_memory()
    rows, mhz = [], []
    for i in range(300):
        env.step_host(act, obs, scans, done)
        out = (ctypes.c_ulonglong * 16)()
        assert lib.f110_debug_phases(out) == 0
        if i >= 100:
            t, c = np.array(list(out)[:8], dtype=np.float64), np.array(list(out)[8:], dtype=np.float64)
            rows.append(np.diff(c[order]))
            mhz.append((c[6] - c[0]) / max(t[6] - t[0], 1.0) * 1e3)
    cyc, f = np.median(np.array(rows), axis=0), float(np.median(mhz))
    print("agents %d: %.1f us inside the kernel at %.0f MHz" % (A, cyc.sum() / f, f))
    for k, v in zip(names, cyc):
        print("    %-56s %7d cycles %6.2f us" % (k, v, v / f))
    env.close()
