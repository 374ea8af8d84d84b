#!/usr/bin/env python
"""Would grouping beams of similar length into the same lockstep round pay?  (CPU only, numpy.)
Replays trace_ray (laser_models.py:107-146) on raceline poses of the bench workload, keeps every beam's visited cells, and
evaluates round schedules: the kernel's (32 consecutive beams) against rounds formed inside windows of W consecutive beams
after sorting the window's beams by their step count -- the count of THIS scan (an upper bound no kernel can have) or of the
scan one tick earlier (what a kernel could keep per beam: a car moves centimetres per tick).  Reports warp iterations per
scan, lane efficiency, and 32-byte sectors per warp-level gather (beams of one round that are no longer neighbours touch more
sectors: the L1 data pipe is 57 % busy in k_raycast32).         python tools/round_grouping_sim.py [n_poses]"""
import os
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from red_gym_b200 import host      # map loading only

mm = host.MapData(os.path.join(host.MAPS_DIR, 'example_map.yaml'), '.png', edt='scipy')
dt, res, ox, oy = mm.edt, mm.resolution, mm.origin[0], mm.origin[1]
H, W_ = dt.shape
wp = np.loadtxt(os.path.join(host.MAPS_DIR, 'example_waypoints.csv'), delimiter=';', skiprows=3)
B, FOV, KMAX = 1080, 2 * np.pi, 256


def march(px, py, yaw):
    """-> steps (n, B) int, rows / cols (n, B, KMAX) int16 of the cell of lookup k = 1 .. steps (lookup 0 is the origin)"""
    n = px.shape[0]
    ang = yaw[:, None] - FOV / 2 + np.arange(B)[None, :] * (FOV / (B - 1))
    cs, sn = np.cos(ang), np.sin(ang)
    x, y = np.repeat(px[:, None], B, 1), np.repeat(py[:, None], B, 1)

    def look(x, y):
        c = np.floor((x - ox) / res).astype(int)
        r = np.floor((y - oy) / res).astype(int)
        oob = (c < 0) | (c >= W_) | (r < 0) | (r >= H)
        v = dt[np.clip(r, 0, H - 1), np.clip(c, 0, W_ - 1)]
        return np.where(oob, dt[-1, -1], v), r, c
    d, _, _ = look(x, y)
    tot = d.copy()
    steps = np.zeros((n, B), int)
    rows = np.zeros((n, B, KMAX), np.int16)
    cols = np.zeros((n, B, KMAX), np.int16)
    alive = (d > 1e-4) & (tot <= 30)
    k = 0
    while alive.any() and k < KMAX:
        x = np.where(alive, x + d * cs, x)
        y = np.where(alive, y + d * sn, y)
        dn, r, c = look(x, y)
        rows[:, :, k] = np.where(alive, np.clip(r, -1, 32000), 0)
        cols[:, :, k] = np.where(alive, np.clip(c, -1, 32000), 0)
        d = np.where(alive, dn, d)
        tot = np.where(alive, tot + dn, tot)
        steps += alive
        alive = alive & (d > 1e-4) & (tot <= 30)
        k += 1
    return steps, rows, cols


def evaluate(groups_of, steps, rows, cols):
    """groups_of(scan index) -> list of beam-id arrays (<= 32 each).  -> iterations / scan, lane efficiency, sectors / gather"""
    it = sec = 0
    for s in range(steps.shape[0]):
        for g in groups_of(s):
            st = steps[s, g]
            m = int(st.max())
            it += m
            for k in range(min(m, KMAX)):
                act = st > k
                sec += np.unique(rows[s, g[act], k].astype(np.int64) * 8192 + (cols[s, g[act], k] >> 3)).size
    n = steps.shape[0]
    return it / n, steps.sum() / 32.0 / it, sec / it


def windows_sorted(key, W):
    def f(s):
        out = []
        for w0 in range(0, B, W):
            ids = np.arange(w0, min(w0 + W, B))
            ids = ids[np.argsort(key[s, ids], kind='stable')]
            out += [ids[i:i + 32] for i in range(0, ids.size, 32)]
        return out
    return f


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    rng = np.random.default_rng(0)
    idx = rng.integers(0, wp.shape[0], n)
    px = wp[idx, 1] + rng.uniform(-0.5, 0.5, n)
    py = wp[idx, 2] + rng.uniform(-0.5, 0.5, n)
    yaw = wp[idx, 3] + np.pi / 2 + rng.uniform(-0.5, 0.5, n)
    steps, rows, cols = march(px, py, yaw)
    # one tick earlier: 5 m/s * 0.01 s back along the heading, yaw rate up to 1 rad/s
    prev, _, _ = march(px - 0.05 * np.cos(yaw), py - 0.05 * np.sin(yaw), yaw - rng.uniform(-0.01, 0.01, n))
    print("%d scans, %.0f lookups per scan, same step count as one tick earlier for %.1f %% of the beams (|diff| <= 1: %.1f %%)"
          % (n, steps.sum(1).mean(), 100 * (steps == prev).mean(), 100 * (np.abs(steps - prev) <= 1).mean()))
    base = evaluate(lambda s: [np.arange(i, min(i + 32, B)) for i in range(0, B, 32)], steps, rows, cols)
    print("rounds of 32 consecutive beams (the kernel):  %.1f iterations / scan, lane efficiency %.3f, %.2f sectors / gather" % base)
    for W in (64, 128, 256, 1080):
        for name, key in (("this scan's counts", steps), ("previous tick's counts", prev)):
            r = evaluate(windows_sorted(key, W), steps, rows, cols)
            print("windows of %4d beams sorted by %-22s: %.1f iterations / scan, lane efficiency %.3f, %.2f sectors / gather  "
                  "(iterations x %.2f, sectors in total x %.2f)" % (W, name, r[0], r[1], r[2], r[0] / base[0], r[0] * r[2] / (base[0] * base[2])))


main()
