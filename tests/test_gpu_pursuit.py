"""GPU parity of the batched pure-pursuit planner (csrc/f110_pursuit.cu through the C ABI) against the unmodified
reference planner's outputs (tests/golden/pursuit.npz, waypoint_follow.npz) and the closed loop of BASELINE config 1."""
import os
from argparse import Namespace

import numpy as np
import pytest

from conftest import GOLDEN, MAPS

pytestmark = pytest.mark.gpu
WB = 0.17145 + 0.15875
CONF = Namespace(wpt_path=os.path.join(MAPS, "example_waypoints.csv"), wpt_delim=";", wpt_rowskip=3, wpt_xind=1, wpt_yind=2,
                 wpt_thind=3, wpt_vind=5)
TLAD, VGAIN = 0.82461887897713965, 1.375


def test_random_poses_match_reference_planner():
    import torch
    from red_gym_b200 import BatchPurePursuitPlanner
    d = np.load(os.path.join(GOLDEN, "pursuit.npz"))
    pl = BatchPurePursuitPlanner(CONF, WB)
    sp, st = pl.plan(d["x"], d["y"], d["theta"], torch.from_numpy(d["lookahead"]), torch.from_numpy(d["vgain"]))
    want = d["speed_steer"]
    assert (sp.cpu().numpy() == want[:, 0]).all()                       # which waypoint: bit-exact
    assert np.abs(st.cpu().numpy() - want[:, 1]).max() <= 1e-12         # CUDA sin/cos/atan vs glibc


def test_nearest_segment_search_prunes_nothing_it_needs():
    """The planner skips groups of segments that cannot hold the nearest point: the nearest index (np.argmin, first
    minimum) must still be the oracle's for cars on the line, between two parts of it, far outside the map, and exactly
    on way points (ties between the two segments that share one)."""
    import pursuit_oracle as po
    from red_gym_b200 import BatchPurePursuitPlanner
    wp = po.load_waypoints(CONF.wpt_path, CONF.wpt_delim, CONF.wpt_rowskip)
    xy = wp[:, [CONF.wpt_xind, CONF.wpt_yind]]
    rng = np.random.default_rng(7)
    lo, hi = xy.min(axis=0), xy.max(axis=0)
    pts = [xy[rng.integers(0, len(xy), 1500)] + rng.normal(0, 0.4, (1500, 2)),            # around the line
           rng.uniform(lo - 5, hi + 5, (1500, 2)),                                          # anywhere over the map
           rng.uniform(-500, 500, (300, 2)),                                                # far away
           xy[rng.integers(0, len(xy), 300)],                                               # on way points
           0.5 * (xy[:-1] + xy[1:])[rng.integers(0, len(xy) - 1, 300)]]                     # on segment mid points
    pts = np.concatenate(pts)
    th = rng.uniform(-np.pi, np.pi, len(pts))
    pl = BatchPurePursuitPlanner(CONF, WB)
    sp, st, picked = pl.plan(pts[:, 0], pts[:, 1], th, 1.5, 1.0, return_picked=True)
    picked = picked.cpu().numpy()
    for k in range(len(pts)):
        (w_sp, w_st), (i, i2) = po.plan(wp, pts[k, 0], pts[k, 1], th[k], 1.5, 1.0, WB, CONF.wpt_xind, CONF.wpt_yind, CONF.wpt_vind,
                                        return_index=True)
        # second index: the intersected segment; the kernel reports the nearest index again when it re-acquires the line
        assert picked[k, 0] == i and picked[k, 1] in ((i2,) if i2 != -2 else (-2, i)), (k, pts[k], picked[k], i, i2)
        assert sp[k].item() == w_sp and abs(st[k].item() - w_st) <= 1e-12


def test_waypoint_follow_pose_action_pairs():
    from red_gym_b200 import BatchPurePursuitPlanner
    d = np.load(os.path.join(GOLDEN, "waypoint_follow.npz"))
    pl = BatchPurePursuitPlanner(CONF, WB)
    poses, acts = d["poses"][:-1], d["actions"]
    sp, st = pl.plan(poses[:, 0], poses[:, 1], poses[:, 2], TLAD, VGAIN)
    assert (sp.cpu().numpy() == acts[:, 1]).all()
    assert np.abs(st.cpu().numpy() - acts[:, 0]).max() <= 1e-12


def test_closed_loop_lap_on_device():
    """BASELINE config 1 with the planner on the device: step and plan alternate without a host round trip; the lap
    must finish like the reference's run (two laps, same step count within the libm tolerance of the steering angle)."""
    from red_gym_b200 import BatchF110Env, BatchPurePursuitPlanner
    d = np.load(os.path.join(GOLDEN, "waypoint_follow.npz"))
    N = 4
    env = BatchF110Env(num_envs=N, num_agents=1, map=os.path.join(MAPS, "example_map"), timestep=0.01, integrator="RK4")
    pl = BatchPurePursuitPlanner(CONF, WB)
    env.reset(np.tile(d["start"], (N, 1, 1)))
    T = d["actions"].shape[0]
    steps = 0
    while steps < T + 50:
        act = pl.plan_actions(env, TLAD, VGAIN)
        _, _, done, _ = env.step(act)
        steps += 1
        if steps % 64 == 0 and bool(done.all()):
            break
    # the same trajectory as the reference at the stated tolerance, sampled along the lap
    ref_pose = d["poses"]
    st = env.state.cpu().numpy()
    assert bool(env.done_bool.all())
    assert (env.lap_counts.cpu().numpy() == 2).all() and not env.collisions.cpu().numpy().any()
    assert abs(float(env.lap_times.cpu().numpy()[0, 0]) - float(d["lap_times"][-1])) <= 0.011
    # replay to the reference's step count and compare poses
    env.reset(np.tile(d["start"], (N, 1, 1)))
    for t in range(1000):
        env.step(pl.plan_actions(env, TLAD, VGAIN))
    st = env.state.cpu().numpy()
    assert np.abs(st[0, :, 0] - ref_pose[1000, 0]).max() < 1e-6 and np.abs(st[1, :, 0] - ref_pose[1000, 1]).max() < 1e-6
