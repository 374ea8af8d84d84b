"""Golden vectors of the pure-pursuit planner: the UNMODIFIED reference planner (examples/waypoint_follow.py,
PurePursuitPlanner.plan and its njit helpers) run in THIS container on random poses -> tests/golden/pursuit.npz.
Test infrastructure only.     python oracle/gen_pursuit_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REF = "/root/reference"


def main():
    sys.path.insert(0, HERE)
    import ref_shim
    ref_shim.install()
    sys.path.insert(0, REF + "/examples")
    import yaml
    from argparse import Namespace
    import waypoint_follow as wf
    os.chdir(REF + "/examples")
    with open("config_example_map.yaml") as f:
        conf = Namespace(**yaml.load(f, Loader=yaml.FullLoader))
    wb = 0.17145 + 0.15875
    planner = wf.PurePursuitPlanner(conf, wb)
    wp = planner.waypoints
    rng = np.random.default_rng(99)
    n = 6000
    k = rng.integers(0, wp.shape[0], n)
    # mostly near the raceline, some 2-20 m away (re-acquire branch), some beyond 20 m (no waypoint: 4.0, 0.0)
    off = np.where(rng.uniform(size=n) < 0.7, rng.normal(0, 0.4, n), rng.uniform(-35, 35, n))
    ang = rng.uniform(0, 2 * np.pi, n)
    x = wp[k, conf.wpt_xind] + off * np.cos(ang)
    y = wp[k, conf.wpt_yind] + off * np.sin(ang)
    th = np.where(rng.uniform(size=n) < 0.5, wp[k, conf.wpt_thind] + np.pi / 2 + rng.normal(0, 0.3, n), rng.uniform(-7, 7, n))
    ld = np.where(rng.uniform(size=n) < 0.6, 0.82461887897713965, rng.uniform(0.2, 5.0, n))       # tlad range of the config
    vg = np.where(rng.uniform(size=n) < 0.6, 1.375, rng.uniform(0.5, 1.5, n))
    x[:3] = wp[[0, 10, 782], conf.wpt_xind]; y[:3] = wp[[0, 10, 782], conf.wpt_yind]          # exactly on waypoints
    out = np.empty((n, 2))
    for j in range(n):
        out[j] = planner.plan(x[j], y[j], th[j], ld[j], vg[j])
    np.savez_compressed(os.path.join(REPO, "tests", "golden", "pursuit.npz"), x=x, y=y, theta=th, lookahead=ld, vgain=vg,
                        speed_steer=out, wheelbase=wb)
    print("wrote pursuit.npz", n, "fallback (4, 0):", int(((out[:, 0] == 4.0) & (out[:, 1] == 0.0)).sum()))


if __name__ == "__main__":
    main()
