#!/bin/bash
# single-env latency with 1 / 2 / 4 agents: f110_step_host per call and F110Env.step of the drop-in
mkdir -p gpurun_out
python - <<'PY' 2>&1 | tail -12
import os, sys, time, ctypes
import numpy as np, torch
sys.path.insert(0, os.getcwd())
from red_gym_b200 import BatchF110Env, MAPS_DIR
import f110_gym
def per_call(f, K=2000, W=300):
    for _ in range(W): f()
    t0 = time.perf_counter()
    for _ in range(K): f()
    return (time.perf_counter() - t0) / K * 1e6
for A in (1, 2, 4):
    env = BatchF110Env(num_envs=1, num_agents=A, map=os.path.join(MAPS_DIR, "example_map"), dtype="fp64", noise="numpy")
    poses = np.array([[[0.7 + 1.2 * a, 0.0, 1.37079632679] for a in range(A)]])
    env.reset(poses); env.reserve_noise(20000)
    act = torch.zeros((1, A, 2), dtype=torch.float64).pin_memory(); act[..., 1] = 1.0
    obs = torch.zeros((10, 1, A), dtype=torch.float64).pin_memory()
    scans = torch.zeros((1, A, 1080), dtype=torch.float64).pin_memory()
    done = torch.zeros((1,), dtype=torch.uint8).pin_memory()
    ptrs = (act.data_ptr(), obs.data_ptr(), scans.data_ptr(), done.data_ptr())
    l0 = env.launch_count
    w = per_call(lambda: env.step_host_ptrs(*ptrs))
    print("agents %d: f110_step_host %.1f us per call, %.1f launches per step" % (A, w, (env.launch_count - l0) / 2300.0))
    env.close()
    genv = f110_gym.make("f110_gym:f110-v0", map=os.path.join(MAPS_DIR, "example_map"), map_ext=".png", num_agents=A)
    genv.reset(poses[0])
    action = np.array([[0.0, 1.0]] * A)
    print("agents %d: F110Env.step %.1f us" % (A, per_call(lambda: genv.step(action), 1500, 200)))
PY
