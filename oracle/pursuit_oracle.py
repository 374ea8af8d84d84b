"""CPU oracle of the pure-pursuit planner.  TEST INFRASTRUCTURE ONLY.

Restates examples/waypoint_follow.py of the reference in plain numpy / Python floats, same operations in the same
order: nearest_point_on_trajectory (:15-46), first_point_on_trajectory_intersecting_circle (:48-127), get_actuation
(:129-141), PurePursuitPlanner._get_current_waypoint / plan (:183-217).  The reference's 2-element np.dot goes through
BLAS inside Numba; here it is a0*b0 + a1*b1.  Pinned (tests/test_pursuit_oracle.py) against tests/golden/pursuit.npz
(random poses through the unmodified reference planner, oracle/gen_pursuit_golden.py) and against the 3329
pose -> action pairs of the reference's own waypoint_follow run (tests/golden/waypoint_follow.npz).
"""
import math

import numpy as np


def load_waypoints(path, delim=";", rowskip=3):
    return np.loadtxt(path, delimiter=delim, skiprows=rowskip)


def nearest_point(point, traj):
    """-> (dist, t, i) of nearest_point_on_trajectory."""
    diffs = traj[1:] - traj[:-1]
    l2s = diffs[:, 0] * diffs[:, 0] + diffs[:, 1] * diffs[:, 1]
    rel = point[None, :] - traj[:-1]
    dots = rel[:, 0] * diffs[:, 0] + rel[:, 1] * diffs[:, 1]
    t = dots / l2s
    t[t < 0.0] = 0.0
    t[t > 1.0] = 1.0
    proj = traj[:-1] + (t * diffs.T).T
    tmp = point[None, :] - proj
    dists = np.sqrt(tmp[:, 0] * tmp[:, 0] + tmp[:, 1] * tmp[:, 1])
    i = int(np.argmin(dists))
    return float(dists[i]), float(t[i]), i


def _dot(a, b):
    return a[0] * b[0] + a[1] * b[1]


def first_point_on_circle(point, radius, traj, t=0.0, wrap=False):
    """-> (first_i, first_t) or (None, None)."""
    n = traj.shape[0]
    start_i = int(t)
    start_t = t % 1.0
    px, py = float(point[0]), float(point[1])

    def seg(i0, i1):
        s = (float(traj[i0, 0]), float(traj[i0, 1]))
        e = (float(traj[i1, 0]) + 1e-6, float(traj[i1, 1]) + 1e-6)
        V = (e[0] - s[0], e[1] - s[1])
        a = _dot(V, V)
        b = 2.0 * _dot(V, (s[0] - px, s[1] - py))
        c = _dot(s, s) + _dot((px, py), (px, py)) - 2.0 * _dot(s, (px, py)) - radius * radius
        disc = b * b - 4 * a * c
        if disc < 0:
            return None
        disc = math.sqrt(disc)
        return (-b - disc) / (2.0 * a), (-b + disc) / (2.0 * a)

    for i in range(start_i, n - 1):
        r = seg(i, i + 1)
        if r is None:
            continue
        t1, t2 = r
        if i == start_i:
            if 0.0 <= t1 <= 1.0 and t1 >= start_t:
                return i, t1
            if 0.0 <= t2 <= 1.0 and t2 >= start_t:
                return i, t2
        elif 0.0 <= t1 <= 1.0:
            return i, t1
        elif 0.0 <= t2 <= 1.0:
            return i, t2
    if wrap:
        for i in range(-1, start_i):
            r = seg(i % n, (i + 1) % n)
            if r is None:
                continue
            t1, t2 = r
            if 0.0 <= t1 <= 1.0:
                return i, t1
            elif 0.0 <= t2 <= 1.0:
                return i, t2
    return None, None


def plan(waypoints, pose_x, pose_y, pose_theta, lookahead_distance, vgain, wheelbase, xind=1, yind=2, vind=5,
         max_reacquire=20.0, return_index=False):
    """PurePursuitPlanner.plan -> (speed, steering_angle) [, (i, i2)]."""
    wpts = np.ascontiguousarray(np.vstack((waypoints[:, xind], waypoints[:, yind])).T)
    pos = np.array([pose_x, pose_y], dtype=np.float64)
    dist, t, i = nearest_point(pos, wpts)
    i2 = None
    if dist < lookahead_distance:
        i2, _ = first_point_on_circle(pos, lookahead_distance, wpts, i + t, wrap=True)
        if i2 is None:
            la = None
        else:
            la = (float(wpts[i2, 0]), float(wpts[i2, 1]), float(waypoints[i, vind]))
    elif dist < max_reacquire:
        la = (float(wpts[i, 0]), float(wpts[i, 1]), float(waypoints[i, vind]))
    else:
        la = None
    if la is None:
        out = (4.0, 0.0)
    else:
        wy = math.sin(-pose_theta) * (la[0] - pose_x) + math.cos(-pose_theta) * (la[1] - pose_y)
        if abs(wy) < 1e-6:
            out = (vgain * la[2], 0.0)
        else:
            radius = 1 / (2.0 * wy / lookahead_distance ** 2)
            out = (vgain * la[2], math.atan(wheelbase / radius))
    if return_index:
        return out, (i, -2 if i2 is None else i2)
    return out
