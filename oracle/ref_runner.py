"""Times the UNMODIFIED reference (F110Env of WE-Autopilot/red_gym's f110_gym: Numba + NumPy) on the host cores.
TEST / MEASUREMENT INFRASTRUCTURE, NOT PRODUCT: used by bench.py's `--impl reference` arm and `cpu_baseline` leg only.

The reference is single-threaded and keeps class-level statics (base_classes.py:63-66), so the natural parallel mode is
one process per core, each with its own F110Env (SURVEY 8d "CPU baseline timing").  The package is
imported from /root/reference/gym (build container) or from oracle/_ref (GPU box; oracle/install_ref.py) through
oracle/ref_shim.py, which only supplies inert `gym` / `pyglet` modules.  This module imports nothing of the product.
"""
import multiprocessing as mp
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def available():
    """(ok, why): the reference can be imported and Numba is there."""
    sys.path.insert(0, HERE)
    try:
        import ref_shim
    except Exception as e:          # pragma: no cover
        return False, "oracle/ref_shim.py not importable: %s" % e
    if not ref_shim.available():
        return False, "reference not installed (oracle/_ref missing: run oracle/install_ref.py where /root/reference exists)"
    try:
        import numba  # noqa: F401
        import scipy  # noqa: F401
        import yaml   # noqa: F401
        import PIL    # noqa: F401
    except Exception as e:
        return False, "a dependency of the reference is missing: %s" % e
    return True, ref_shim.package_parent()


def _poses(csv_path, n, agents, seed=4):
    """Start poses of the bench workload (bench.make_poses): k * R / n along the raceline; further agents 1.5 m behind."""
    wp = np.loadtxt(csv_path, delimiter=";", skiprows=3)
    idx = (np.arange(n) * wp.shape[0]) // n
    lead = np.stack([wp[idx, 1], wp[idx, 2], wp[idx, 3] + np.pi / 2], axis=1)
    rng = np.random.default_rng(seed)
    out = [lead]
    for a in range(1, agents):
        lat = rng.uniform(-0.3, 0.3, n)
        c = lead.copy()
        c[:, 0] += -1.5 * a * np.cos(lead[:, 2]) - lat * np.sin(lead[:, 2])
        c[:, 1] += -1.5 * a * np.sin(lead[:, 2]) + lat * np.cos(lead[:, 2])
        out.append(c)
    return np.stack(out, axis=1)


def _worker(rank, world, agents, map_base, csv_path, seconds, solo_seconds, ready, go_solo, go_all, out):
    os.environ.setdefault("NUMBA_CACHE_DIR", "/tmp/f110_ref_numba_cache")
    os.environ["NUMBA_NUM_THREADS"] = "1"
    try:
        os.sched_setaffinity(0, {sorted(os.sched_getaffinity(0))[rank % len(os.sched_getaffinity(0))]})
    except Exception:
        pass
    sys.path.insert(0, HERE)
    import ref_shim
    gym = ref_shim.install()
    from f110_gym.envs.base_classes import Integrator
    env = gym.make("f110_gym:f110-v0", map=map_base, map_ext=".png", num_agents=agents, timestep=0.01, integrator=Integrator.RK4, seed=12345)
    pool = _poses(csv_path, 1024, agents)
    rng = np.random.default_rng(2024 + rank)
    k = (rank * 1024) // max(world, 1)

    def act():
        return np.stack([rng.uniform(-0.4189, 0.4189, agents), rng.uniform(0.0, 8.0, agents)], axis=-1)

    env.reset(pool[k % 1024].copy())
    for _ in range(200):            # warm-up 200 steps (JIT compilation, caches)
        obs, r, done, info = env.step(act())
        if done:
            k += 1
            env.reset(pool[k % 1024].copy())

    def run(limit):
        nonlocal k
        steps, t0 = 0, time.perf_counter()
        while True:
            obs, r, done, info = env.step(act())
            steps += 1
            if done:                # auto-reset like the GPU workload (a reset = set poses + one zero-action step)
                k += 1
                env.reset(pool[k % 1024].copy())
                steps += 1
            if (steps & 15) == 0 and time.perf_counter() - t0 >= limit:
                break
        return steps, time.perf_counter() - t0

    ready.wait()
    solo = None
    if rank == 0 and solo_seconds > 0:
        solo = run(solo_seconds)
    go_solo.wait()
    both = run(seconds)
    go_all.wait()
    out.put((rank, solo, both))


def run(agents=1, seconds=12.0, solo_seconds=4.0, procs=0, maps_dir=None):
    """One F110Env per process on `procs` cores (default: all the process may use).  Returns a dict with the aggregate
    agent-steps/s of all processes running at once, the single-process figure, core count and CPU model."""
    maps_dir = maps_dir or os.path.join(os.path.dirname(HERE), "red_gym_b200", "maps")      # same example_map.png/.yaml bytes as the reference's examples/
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    procs = procs or int(os.environ.get("F110_REF_PROCS", "0")) or cores
    ctx = mp.get_context("spawn")
    ready, go_solo, go_all = ctx.Barrier(procs), ctx.Barrier(procs), ctx.Barrier(procs)
    q = ctx.Queue()
    t_start = time.perf_counter()
    ps = [ctx.Process(target=_worker, args=(r, procs, agents, os.path.join(maps_dir, "example_map"), os.path.join(maps_dir, "example_waypoints.csv"),
                                            seconds, solo_seconds, ready, go_solo, go_all, q), daemon=True) for r in range(procs)]
    for p in ps:
        p.start()
    res = []
    deadline = time.time() + 600 + seconds + solo_seconds
    while len(res) < procs and time.time() < deadline:
        try:
            res.append(q.get(timeout=5.0))
        except Exception:
            if any(p.exitcode not in (None, 0) for p in ps):
                break
    for p in ps:
        p.join(timeout=5.0)
        if p.is_alive():
            p.terminate()
    if len(res) < procs:
        raise RuntimeError("reference workers failed (%d of %d reported)" % (len(res), procs))
    solo = [s for _, s, _ in res if s is not None]
    total = sum(b[0] for _, _, b in res) * agents
    wall = max(b[1] for _, _, b in res)
    return {"value": total / wall, "unit": "agent-steps/s", "cores": procs, "kind": "reference",
            "single_process_value": (solo[0][0] * agents / solo[0][1]) if solo else None,
            "cpu_model": cpu_model(), "host_cpus": os.cpu_count(), "timed_seconds": wall, "total_seconds": time.perf_counter() - t_start,
            "sample_steps": int(sum(b[0] for _, _, b in res)),
            "sample": "unmodified reference F110Env (Numba/NumPy, oracle/_ref), one process per core x %d, %d agent(s), example_map, RK4, 1080 beams, "
                      "numpy noise, random actions, reset on done; 200 warm-up steps per process, then %.1f s timed" % (procs, agents, wall)}


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.lower().startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


if __name__ == "__main__":
    import json
    print(json.dumps(run(agents=int(sys.argv[1]) if len(sys.argv) > 1 else 1, seconds=float(sys.argv[2]) if len(sys.argv) > 2 else 5.0)))
