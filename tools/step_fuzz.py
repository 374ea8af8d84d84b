#!/usr/bin/env python
"""Random step configurations against the oracle: 1 .. 6 agents per env, 2 .. 1300 beams, fields of view 0.5 .. 2 pi, RK4 / Euler,
time steps 2 .. 20 ms, four maps, perturbed vehicle parameters, ego index anywhere, random actions, cars started close to each other
so that GJK, the opponent casts and iTTC all fire.  fp64 parity path (fused / split / whole-scan kernels by size): flags, counters and
collision partners bit-exact, states to 1e-9 relative, scans to 1e-4 m, an env that diverged after a discontinuity is counted.
With F110_B200_LIB=red_gym_b200/libf110_b200_checked.so the bounds-checked build runs and its counters are printed at the end.
    python tools/step_fuzz.py [configs=30] [seed=1] [only=config number: the others are skipped, their random draws are not]"""
import ctypes, os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "oracle"))
import f110_oracle as fo
from red_gym_b200 import BatchF110Env, MAPS_DIR, _native
from red_gym_b200.host import raceline_poses

K = int(sys.argv[1]) if len(sys.argv) > 1 else 30
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
ONLY = int(sys.argv[3]) if len(sys.argv) > 3 else -1
REL, SCAN_TOL = 1e-9, 1e-4
findings = 0
for k in range(K):
    A = int(rng.choice([1, 1, 2, 2, 3, 4, 6]))
    N = int(rng.choice([1, 7, 60, 300, 700]))
    B = int(rng.choice([2, 3, 17, 32, 33, 100, 256, 540, 1080, int(rng.integers(2, 1300))]))
    fov = float(rng.choice([4.7, 2 * np.pi, float(rng.uniform(0.5, 2 * np.pi))]))
    integ = str(rng.choice(["RK4", "Euler"]))
    dt = float(rng.choice([0.01, 0.002, 0.02]))
    m = str(rng.choice(["example_map", "berlin", "vegas", "skirk"]))
    T = int(rng.choice([20, 60, 120]))
    ego = int(rng.integers(0, A))
    params = dict(fo.DEFAULT_PARAMS)
    if rng.random() < 0.5:
        for key in ("mu", "C_Sf", "C_Sr", "m", "I", "a_max", "v_max", "sv_max"):
            params[key] = params[key] * float(rng.uniform(0.8, 1.25))
        params["sv_min"] = -params["sv_max"]
    if m == "example_map":
        base = raceline_poses(N)
    else:      # the other maps have no raceline here: near their origin, where the track is
        base = np.stack([rng.uniform(-1.0, 1.0, N), rng.uniform(-1.0, 1.0, N), rng.uniform(-3, 3, N)], axis=1)
    poses = np.repeat(base[:, None, :], A, axis=1)
    if A > 1:
        poses[:, 1:, 0] += rng.uniform(-2.0, 2.0, (N, A - 1))
        poses[:, 1:, 1] += rng.uniform(-2.0, 2.0, (N, A - 1))
        poses[:, 1:, 2] = rng.uniform(-3, 3, (N, A - 1))
    kw = dict(map=os.path.join(MAPS_DIR, m), seed=int(rng.integers(1, 1 << 30)), ego_idx=ego, num_beams=B, fov=fov, integrator=integ,
              timestep=dt, params=params)
    if ONLY >= 0 and k != ONLY:
        for t in range(T):
            rng.uniform(-0.4, 0.4, (N, A)); rng.uniform(0.0, 8.0, (N, A))
        continue
    env = BatchF110Env(num_envs=N, num_agents=A, **kw)
    ora = fo.OracleEnv(N, A, noise_rows=T + 8, **kw)
    env.reset(poses); ora.reset(poses)
    bad = np.zeros(N, dtype=bool)
    first = {}
    hits = ttc = 0
    for t in range(T + 1):
        if t > 0:
            act = np.stack([rng.uniform(-0.4, 0.4, (N, A)), rng.uniform(0.0, 8.0, (N, A))], axis=-1)
            env.step(act); ora.step(act)
        st = env.state.permute(1, 2, 0).cpu().numpy()
        checks = {
            "collisions": (env.collisions.cpu().numpy() == ora.collisions).all(axis=1),
            "collision_idx": (env.collision_idx.cpu().numpy() == ora.collision_idx).all(axis=1),
            "toggles": (env.toggle_list.cpu().numpy() == ora.toggles).all(axis=1),
            "done": env.done.cpu().numpy().astype(bool) == ora.done.astype(bool),
            "state": (np.abs(st - ora.state) <= REL * np.maximum(1.0, np.abs(ora.state))).all(axis=(1, 2)),
            "scans": (np.abs(env.scans.cpu().numpy() - ora.scans) > SCAN_TOL).mean(axis=(1, 2)) < (1e-3 if B >= 100 else 1.0 / (A * B) - 1e-12),
        }
        newly = np.zeros(N, dtype=bool)
        for name, ok in checks.items():
            new = ~ok & ~bad
            newly |= new
            for e in np.flatnonzero(new):
                first.setdefault(int(e), (t, []))[1].append(name)
        for e in np.flatnonzero(newly):      # what the first difference of an env looks like
            ds = np.abs(env.scans[e].cpu().numpy() - ora.scans[e])
            if ONLY >= 0:
                np.set_printoptions(precision=17, linewidth=200)
                print("env %d step %d: oracle states (x, y, delta, v, yaw, yaw rate, slip) per agent:" % (e, t)); print(ora.state[e])
                print("state difference:"); print(st[e] - ora.state[e])
                for a in range(A):
                    off = np.flatnonzero(ds[a] > SCAN_TOL)
                    if off.size:
                        print("agent %d: beams off %s" % (a, off))
                        print("   library:", env.scans[e, a].cpu().numpy()[off][:12]); print("   oracle: ", ora.scans[e, a][off][:12])
            first[int(e)] = first[int(e)] + ("beams off %d, largest %.3g m" % (int((ds > SCAN_TOL).sum()), float(ds.max())),
                                             "state diff %.3g" % float(np.abs(st[e] - ora.state[e]).max()))
        bad |= newly
        hits += int((ora.collision_idx >= 0).sum()); ttc += int(ora.collisions.sum())
    # a beam grazing a corner may fall on the other side in the last ulp of sin / cos; everything after it then differs
    tol = max(1, N // 60) if A > 1 else max(0, N // 300)
    msg = ("ok" if bad.sum() <= tol else "DIVERGED envs %d (allowed %d)" % (int(bad.sum()), tol)) + (" first differences (env: step, what) %s" % first if first else "")
    findings += bad.sum() > tol
    print("config %2d: A %d N %3d B %4d fov %.3f %-5s dt %.3f %-11s T %3d ego %d  gjk %5d coll %5d  diverged %d  %s"
          % (k, A, N, B, fov, integ, dt, m, T, ego, hits, ttc, int(bad.sum()), msg), flush=True)
    env.close()
L = _native.load()
out = (ctypes.c_uint64 * 8)()
L.f110_debug_violations.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int32]
L.f110_debug_violations(0, out, 0)
print("checked build: %s; violation counters %s; configs with a finding: %d of %d" % (bool(L.f110_build_is_checked()), list(out), findings, K))
sys.exit(1 if findings or any(list(out)[:7]) else 0)
