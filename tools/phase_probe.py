#!/usr/bin/env python
"""Where the time of a single-env fused step goes: clock64 stamps taken by thread 0 of env 0 inside k_step_fused64
(diagnostic build: python red_gym_b200/build.py --phases).   F110_B200_LIB=red_gym_b200/libf110_b200_phases.so python tools/phase_probe.py"""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from red_gym_b200 import BatchF110Env, MAPS_DIR, _native

lib = _native.load()
lib.f110_debug_phases.argtypes = [ctypes.c_void_p]
names = ["dynamics (RK4, thread 0)", "scan preparation (thread 0)", "barrier + ray-cast rounds of thread 0's warp", "barrier: the other warps' rounds",
         "iTTC outcome, lap bookkeeping, obs export", "scan export to the host (rest of the CTA)"]
# variant 1: 256 threads per env even with fewer envs than SMs (default then: 1024)
for n, variant, with_scans in ((1, 0, True), (1, 0, False), (1, 1, True)):
    env = BatchF110Env(num_envs=n, num_agents=1, map=os.path.join(MAPS_DIR, "example_map"), dtype="fp64", noise="numpy", raycast_variant=variant)
    env.reset(np.tile(np.array([[[0.7, 0.0, 1.37079632679]]]), (n, 1, 1)))
    act = torch.zeros((n, 1, 2), dtype=torch.float64).pin_memory()
    obs = torch.zeros((10, n, 1), dtype=torch.float64).pin_memory()
    scans = torch.zeros((n, 1, 1080), dtype=torch.float64).pin_memory()
    done = torch.zeros((n,), dtype=torch.uint8).pin_memory()
    rows, mhz = [], []
    for i in range(300):
        env.step_host(act, obs, scans if with_scans else None, done)
        out = (ctypes.c_ulonglong * 16)()
        assert lib.f110_debug_phases(out) == 0
        if i >= 100:
            t, c = np.array(list(out)[:8], dtype=np.float64), np.array(list(out)[8:], dtype=np.float64)
            rows.append(np.diff(c[:7]))
            mhz.append((c[6] - c[0]) / max(t[6] - t[0], 1.0) * 1e3)
    cyc, f = np.median(np.array(rows), axis=0), float(np.median(mhz))
    print("envs %d, %s threads per env, scan exported to the host: %s -- %.1f us inside the kernel at %.0f MHz"
          % (n, "256" if variant & 1 else "1024", with_scans, cyc.sum() / f, f))
    for k, v in zip(names, cyc):
        print("    %-48s %7d cycles %6.2f us" % (k, v, v / f))
    env.close()
