#!/usr/bin/env python
"""Host side of a single-env step: wall time per call of (a) an empty ctypes call, (b) f110_step_host called with prebuilt
ctypes arguments, (c) BatchF110Env.step_host_ptrs, (d) BatchF110Env.step_host, (e) the drop-in F110Env.step.
   python tools/host_overhead_probe.py"""
import ctypes, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from red_gym_b200 import BatchF110Env, MAPS_DIR, _native

def per_call(f, K=3000, W=300):
    for _ in range(W):
        f()
    t0 = time.perf_counter()
    for _ in range(K):
        f()
    return (time.perf_counter() - t0) / K * 1e6

lib = _native.load()
print("empty ctypes call (f110_abi_version): %.2f us" % per_call(lib.f110_abi_version, 20000))
for scans_too in (True, False):
    env = BatchF110Env(num_envs=1, num_agents=1, map=os.path.join(MAPS_DIR, "example_map"), dtype="fp64", noise="numpy")
    env.reset(np.array([[[0.7, 0.0, 1.37079632679]]]))
    env.reserve_noise(20000)
    act = torch.zeros((1, 1, 2), dtype=torch.float64).pin_memory()
    obs = torch.zeros((10, 1, 1), dtype=torch.float64).pin_memory()
    scans = torch.zeros((1, 1, 1080), dtype=torch.float64).pin_memory() if scans_too else None
    done = torch.zeros((1,), dtype=torch.uint8).pin_memory()
    vp = ctypes.c_void_p
    args = (env._h, vp(act.data_ptr()), vp(obs.data_ptr()), vp(scans.data_ptr()) if scans_too else None, vp(done.data_ptr()), vp(env._raw_stream(env.device_index) or 0))
    raw = per_call(lambda: lib.f110_step_host(*args))
    ptrs = (act.data_ptr(), obs.data_ptr(), scans.data_ptr() if scans_too else None, done.data_ptr())
    a = per_call(lambda: env.step_host_ptrs(*ptrs))
    b = per_call(lambda: env.step_host(act, obs, scans, done))
    print("scans exported: %s | f110_step_host raw %.1f us | step_host_ptrs %.1f us | step_host %.1f us" % (scans_too, raw, a, b))
    env.close()

import f110_gym
env = f110_gym.make("f110_gym:f110-v0", map=os.path.join(MAPS_DIR, "example_map"), map_ext=".png", num_agents=1)
env.reset(np.array([[0.7, 0.0, 1.37079632679]]))
action = np.array([[0.0, 0.0]])
print("F110Env.step (drop-in, obs dict built): %.1f us" % per_call(lambda: env.step(action), 2000, 200))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(2000):
    env.step(action)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(12)
