#!/usr/bin/env python
"""BASELINE config 1 on the B200 library: one car follows example_waypoints.csv with pure pursuit
(the reference's examples/waypoint_follow.py, same map, start pose, look-ahead and speed gain).

    python examples/waypoint_follow.py                 # the reference's loop: F110Env drop-in, one env, host planner call per step
    python examples/waypoint_follow.py --envs 4096     # the same lap in 4096 envs at once, step and plan both on the device

Prints the simulated lap time and the wall time, like the reference's script.
"""
import argparse
import os
import sys
import time
from argparse import Namespace

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
MAPS = os.path.join(REPO, "red_gym_b200", "maps")
CONF = Namespace(wpt_path=os.path.join(MAPS, "example_waypoints.csv"), wpt_delim=";", wpt_rowskip=3, wpt_xind=1, wpt_yind=2,
                 wpt_thind=3, wpt_vind=5, sx=0.7, sy=0.0, stheta=1.37079632679)      # examples/config_example_map.yaml
TLAD, VGAIN = 0.82461887897713965, 1.375
WHEELBASE = 0.17145 + 0.15875


def single_env():
    import f110_gym
    from red_gym_b200 import BatchPurePursuitPlanner
    env = f110_gym.make("f110_gym:f110-v0", map=os.path.join(MAPS, "example_map"), map_ext=".png", num_agents=1, timestep=0.01)
    planner = BatchPurePursuitPlanner(CONF, WHEELBASE)
    obs, step_reward, done, info = env.reset(np.array([[CONF.sx, CONF.sy, CONF.stheta]]))
    laptime, start = 0.0, time.time()
    while not done:
        speed, steer = planner.plan(obs["poses_x"][0], obs["poses_y"][0], obs["poses_theta"][0], TLAD, VGAIN)
        obs, step_reward, done, info = env.step(np.array([[float(steer[0]), float(speed[0])]]))
        laptime += step_reward
    print("Sim elapsed time:", laptime, "Real elapsed time:", time.time() - start)


def batched(n):
    from red_gym_b200 import BatchF110Env, BatchPurePursuitPlanner
    env = BatchF110Env(num_envs=n, num_agents=1, map=os.path.join(MAPS, "example_map"), timestep=0.01)
    planner = BatchPurePursuitPlanner(CONF, WHEELBASE)
    env.reset(np.tile(np.array([[[CONF.sx, CONF.sy, CONF.stheta]]]), (n, 1, 1)))
    steps, start = 0, time.time()
    while True:
        _, _, done, _ = env.step(planner.plan_actions(env, TLAD, VGAIN))
        steps += 1
        if steps % 64 == 0 and bool(done.all()):      # one device -> host read every 64 steps
            break
    wall = time.time() - start
    print("envs:", n, "steps:", steps, "sim lap time of env 0:", float(env.lap_times[0, 0]), "real elapsed time:", wall,
          "agent-steps/s:", n * steps / wall)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=1)
    a = ap.parse_args()
    single_env() if a.envs == 1 else batched(a.envs)
