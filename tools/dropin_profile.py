#!/usr/bin/env python
"""Where a single-env drop-in step spends its time: python tools/dropin_profile.py [steps]"""
import cProfile, os, pstats, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from f110_gym.envs import F110Env
from f110_gym.envs.base_classes import Integrator
from red_gym_b200 import MAPS_DIR

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
env = F110Env(map=os.path.join(MAPS_DIR, "example_map"), map_ext=".png", num_agents=1, timestep=0.01, integrator=Integrator.RK4)
env.reset(np.array([[0.7, 0.0, 1.37079632679]]))
act = np.array([[0.0, 1.0]])
for _ in range(200):
    env.step(act)
t0 = time.perf_counter()
for _ in range(n):
    env.step(act)
dt = time.perf_counter() - t0
print("F110Env.step: %.1f us per step" % (1e6 * dt / n))
b = env.sim._batch
sim = env.sim
t0 = time.perf_counter()
for _ in range(n):
    b.step_host(sim._h_act, sim._h_obs, sim._h_scans, sim._h_done)
dt = time.perf_counter() - t0
print("BatchF110Env.step_host (C ABI call incl. copies and sync): %.1f us per step" % (1e6 * dt / n))
env.reset(np.array([[0.7, 0.0, 1.37079632679]]))
pr = cProfile.Profile()
pr.enable()
for _ in range(n):
    env.step(act)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(12)
