"""CPU: the pure-pursuit oracle (oracle/pursuit_oracle.py) against the unmodified reference planner: random poses
(tests/golden/pursuit.npz) and the 3329 pose -> action pairs of the reference's waypoint_follow run."""
import os

import numpy as np

import pursuit_oracle as po
from conftest import GOLDEN, MAPS

WP = po.load_waypoints(os.path.join(MAPS, "example_waypoints.csv"))
WB = 0.17145 + 0.15875


def test_random_poses_match_reference_planner():
    d = np.load(os.path.join(GOLDEN, "pursuit.npz"))
    n = d["x"].shape[0]
    got = np.array([po.plan(WP, d["x"][j], d["y"][j], d["theta"][j], d["lookahead"][j], d["vgain"][j], WB) for j in range(n)])
    want = d["speed_steer"]
    assert (got[:, 0] == want[:, 0]).all()                    # speed = vgain * table value: which waypoint was picked
    assert np.abs(got[:, 1] - want[:, 1]).max() <= 1e-12      # steering angle (BLAS dot vs a0*b0 + a1*b1, libm atan)
    assert ((want[:, 0] == 4.0) & (want[:, 1] == 0.0)).sum() > 100      # the no-waypoint branch is covered


def test_waypoint_follow_lap_actions():
    d = np.load(os.path.join(GOLDEN, "waypoint_follow.npz"))
    poses, acts = d["poses"], d["actions"]
    for t in range(0, acts.shape[0], 3):
        sp, st = po.plan(WP, poses[t, 0], poses[t, 1], poses[t, 2], 0.82461887897713965, 1.375, WB)
        assert sp == acts[t, 1] and abs(st - acts[t, 0]) <= 1e-12, t
