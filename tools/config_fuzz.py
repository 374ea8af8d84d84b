#!/usr/bin/env python
"""Random scan configurations against the oracle: beam counts 2 .. 2300, fields of view 0.05 .. 7 rad (the reference takes any),
four maps, straight and rotated origins, poses inside / at the border of / outside the map; fp64 must be bit-exact (both the
CTA-per-scan and the warp-per-scan kernel, a sorted large launch too), fp32 within the documented tolerance.  With
F110_B200_LIB=red_gym_b200/libf110_b200_checked.so the bounds-checked build runs and its counters are printed at the end.
    python tools/config_fuzz.py [configs=40] [seed=1]"""
import ctypes, os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "oracle"))
import f110_oracle as fo
from red_gym_b200 import BatchF110Env, MAPS_DIR, _native

K = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
bad = 0
for k in range(K):
    B = int(rng.choice([2, 3, 5, 17, 31, 32, 33, 64, 100, 255, 256, 257, 540, 1080, 2160, int(rng.integers(2, 2300))]))
    fov = float(rng.choice([0.05, 1.0, 4.7, 2 * np.pi, 6.5, 7.0, float(rng.uniform(0.05, 7.0))]))
    m = str(rng.choice(["example_map", "berlin", "vegas", "skirk"]))
    yaw = None if rng.random() < 0.6 else float(rng.uniform(-3.1, 3.1))
    n = int(rng.choice([7, 300, 1500, 1500, 40000], p=[0.3, 0.3, 0.2, 0.15, 0.05]))       # 40000: the spatially ordered launch
    kw_o = dict(map=os.path.join(MAPS_DIR, m), noise=False, num_beams=B, fov=fov)
    if yaw is not None:
        kw_o["origin_yaw"] = yaw
    ora = fo.OracleEnv(1, 1, **kw_o)
    span = float(rng.choice([0.5, 5.0, 60.0]))
    poses = np.stack([rng.uniform(-span, span, n), rng.uniform(-span, span, n), rng.uniform(-7, 7, n)], axis=1)
    poses[0] = [1e3, -1e3, 0.5]
    want = ora.scan(poses)
    kw = dict(num_envs=1, num_agents=1, map=os.path.join(MAPS_DIR, m), num_beams=B, fov=fov, noise="off")
    if yaw is not None:
        kw["origin_yaw"] = yaw
    msg = []
    for variant in (0, 128):
        got = BatchF110Env(dtype="fp64", raycast_variant=variant, **kw).raycast(poses).cpu().numpy()
        if not (got == want).all():
            msg.append("fp64 variant %d differs: %d ranges, max %.3g" % (variant, int((got != want).sum()), float(np.abs(got - want).max())))
    g32 = [BatchF110Env(dtype="fp32", raycast_variant=v, **kw).raycast(poses).cpu().numpy() for v in (0, 128)]
    if not (g32[0] == g32[1]).all():
        msg.append("fp32 kernels differ from each other")
    d = np.abs(g32[0] - want)
    if not (np.median(d) < 1e-3 and (d > 2e-2).mean() < 0.05):
        msg.append("fp32 off: median %.3g, %.2f %% beyond 2 cm" % (float(np.median(d)), 100 * float((d > 2e-2).mean())))
    bad += bool(msg)
    print("config %2d: B %4d fov %.3f %-11s yaw %-6s n %4d span %4.1f  %s" % (k, B, fov, m, "-" if yaw is None else "%.2f" % yaw, n, span, "; ".join(msg) or "ok"))
L = _native.load()
out = (ctypes.c_uint64 * 8)()
L.f110_debug_violations.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int32]
L.f110_debug_violations(0, out, 0)
print("checked build: %s; violation counters %s; configs with a finding: %d of %d" % (bool(L.f110_build_is_checked()), list(out), bad, K))
sys.exit(1 if bad or any(list(out)[:7]) else 0)
