"""Golden-vector generator: runs the UNMODIFIED reference (via oracle/ref_shim.py) in THIS
container and writes small fixtures to tests/golden/.  Test infrastructure only.

    python oracle/gen_golden.py            # regenerate everything (one subprocess per scenario)
    python oracle/gen_golden.py --only rollout_a1_rk4

Why subprocesses: `RaceCar.scan_simulator/cosines/scan_angles/side_distances` are class-level
statics (reference base_classes.py:63-66,116) so one process can hold one (beams, fov) only.

What each fixture pins (reference file:line):
  scans_*        ScanSimulator2D.scan(pose, None)            laser_models.py:429 -> get_scan :149 -> trace_ray :107
  rollout_*      F110Env.reset/step                          f110_env.py:261,304 -> Simulator.step base_classes.py:546
                 (states, collisions, collision_idx, toggles, lap counters, scans at sampled steps)
  waypoint_follow  examples/waypoint_follow.py main loop (BASELINE config 1), action stream recorded
"""
import argparse
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
GOLDEN = os.path.join(REPO, "tests", "golden")
REF = "/root/reference"
MAPS = {
    "example_map": REF + "/examples/example_map",
    "berlin": REF + "/gym/f110_gym/envs/maps/berlin",
    "skirk": REF + "/gym/f110_gym/envs/maps/skirk",
    "vegas": REF + "/gym/f110_gym/envs/maps/vegas",
    "stata_basement": REF + "/gym/f110_gym/envs/maps/stata_basement",
}
FOVS = {"2pi": 2 * np.pi, "4p7": 4.7}


def _ref():
    sys.path.insert(0, HERE)
    import ref_shim
    return ref_shim.install()


def _rot_tag(yaw):
    return ("rot%+.1f" % yaw).replace("+", "p").replace("-", "m").replace(".", "d")


def rotated_map(map_name, yaw):
    """The bundled map with origin[2] = yaw: a temporary yaml + a copy of the same image (every bundled yaml has
    yaw 0, so the rotation of xy_2_rc, laser_models.py:75-76 with orig_s/orig_c from :421-422, is otherwise never
    exercised).  Returns the path without extension, as F110Env's `map` kwarg wants it."""
    import shutil
    import yaml
    d = "/tmp/f110_rot_maps"
    os.makedirs(d, exist_ok=True)
    with open(MAPS[map_name] + ".yaml") as f:
        meta = yaml.safe_load(f)
    base = os.path.join(d, "%s_%s" % (map_name, _rot_tag(yaw)))
    meta["origin"] = [float(meta["origin"][0]), float(meta["origin"][1]), float(yaw)]
    meta["image"] = os.path.basename(base) + ".png"
    with open(base + ".yaml", "w") as f:
        yaml.safe_dump(meta, f)
    shutil.copyfile(MAPS[map_name] + ".png", base + ".png")
    return base


def to_rotated_world(poses, map_name, yaw):
    """Poses given in the world frame of the unrotated map -> the same places of the track in the world frame of the
    map whose origin is rotated by yaw (world = origin + R(yaw) * (old - origin), heading + yaw)."""
    import yaml
    with open(MAPS[map_name] + ".yaml") as f:
        o = yaml.safe_load(f)["origin"]
    c, s = np.cos(yaw), np.sin(yaw)
    out = np.array(poses, dtype=np.float64)
    dx, dy = out[..., 0] - o[0], out[..., 1] - o[1]
    out[..., 0] = o[0] + dx * c - dy * s
    out[..., 1] = o[1] + dx * s + dy * c
    out[..., 2] = out[..., 2] + yaw
    return out


def waypoint_poses(n, offset=0.0):
    """n start poses on the example_map raceline: rows k*R/n, (x, y, psi+pi/2)."""
    wp = np.loadtxt(REF + "/examples/example_waypoints.csv", delimiter=";", skiprows=3)
    idx = (np.arange(n) * wp.shape[0]) // n
    poses = np.stack([wp[idx, 1], wp[idx, 2], wp[idx, 3] + np.pi / 2], axis=1)
    if offset:
        poses[:, 0] -= offset * np.cos(poses[:, 2])
        poses[:, 1] -= offset * np.sin(poses[:, 2])
    return poses


# ----------------------------------------------------------------------------- scans
def gen_scans(map_name, fov_name, beams=1080, n_poses=20, yaw=0.0):
    _ref()
    from f110_gym.envs.laser_models import ScanSimulator2D
    fov = FOVS[fov_name]
    sim = ScanSimulator2D(beams, fov)
    sim.set_map((rotated_map(map_name, yaw) if yaw else MAPS[map_name]) + ".yaml", ".png")
    rng = np.random.default_rng(7)
    free = np.argwhere(sim.dt > 0.3)
    pick = free[rng.choice(free.shape[0], n_poses, replace=False)]
    # cell -> map frame (yaw 0: also the world frame; rotated below)
    x = sim.orig_x + (pick[:, 1] + rng.uniform(0, 1, n_poses)) * sim.map_resolution
    y = sim.orig_y + (pick[:, 0] + rng.uniform(0, 1, n_poses)) * sim.map_resolution
    th = rng.uniform(-2 * np.pi, 4 * np.pi, n_poses)   # includes negative / >2pi yaw (wrap path)
    if map_name == "example_map":
        # most free cells of example_map lie outside the track: put 3/4 of the poses on the raceline
        wp = np.loadtxt(REF + "/examples/example_waypoints.csv", delimiter=";", skiprows=3)
        k = (3 * n_poses) // 4
        rows = rng.choice(wp.shape[0], k, replace=False)
        x[:k] = wp[rows, 1] + rng.uniform(-0.4, 0.4, k)
        y[:k] = wp[rows, 2] + rng.uniform(-0.4, 0.4, k)
    poses = np.stack([x, y, th], axis=1)
    # edge cases: outside the map (xy_2_rc -> (-1,-1), laser_models.py:79-81), on an obstacle,
    # exactly on the map origin, yaw exactly 0 and fov/2
    occ = np.argwhere(sim.dt == 0.0)[0]
    extra = np.array([
        [sim.orig_x - 5.0, sim.orig_y - 5.0, 0.3],
        [sim.orig_x + sim.map_width * sim.map_resolution + 1.0, sim.orig_y + 1.0, 2.0],
        [sim.orig_x + (occ[1] + 0.5) * sim.map_resolution, sim.orig_y + (occ[0] + 0.5) * sim.map_resolution, 1.0],
        [sim.orig_x, sim.orig_y, 0.0],
        [x[0], y[0], 0.0],
        [x[0], y[0], fov / 2.0],
    ])
    poses = np.concatenate([poses, extra], axis=0)
    if yaw:
        poses = to_rotated_world(poses, map_name, yaw)
        poses[-2:, 2] = [0.0, fov / 2.0]          # keep the two exact-heading cases exact
    scans = np.stack([sim.scan(p, None) for p in poses])
    if yaw:
        out = os.path.join(GOLDEN, "scans_%s_%s_b%d_%s.npz" % (map_name, fov_name, beams, _rot_tag(yaw)))
        np.savez_compressed(out, poses=poses, scans=scans, fov=fov, beams=beams, map=map_name, origin_yaw=yaw)
        assert (scans[:n_poses] < 29.0).any(axis=1).mean() > 0.5      # the poses do sit inside the rotated track
    else:
        out = os.path.join(GOLDEN, "scans_%s_%s_b%d.npz" % (map_name, fov_name, beams))
        np.savez_compressed(out, poses=poses, scans=scans, fov=fov, beams=beams, map=map_name)
    print("wrote", out, scans.shape)


def gen_scans_many_beams(beams=20001):
    """A beam count far above the env's 1080: theta_index_increment ~ 0.1, so the reference's running sum
    (laser_models.py:180-184) passes an integer every ten beams and its accumulated rounding decides int() -- the case
    the product's closed-form beam index has to replay exactly (f110_kernels_fp64.cu: theta_index_of)."""
    _ref()
    from f110_gym.envs.laser_models import ScanSimulator2D
    sim = ScanSimulator2D(beams, 2 * np.pi)
    sim.set_map(MAPS["example_map"] + ".yaml", ".png")
    p = waypoint_poses(3)
    poses = np.array([[p[0, 0], p[0, 1], 0.0], [p[1, 0], p[1, 1], 1.2345678], [p[2, 0], p[2, 1], np.pi]])
    scans = np.stack([sim.scan(q, None) for q in poses])
    out = os.path.join(GOLDEN, "scans_example_map_2pi_b%d.npz" % beams)
    np.savez_compressed(out, poses=poses, scans=scans, fov=2 * np.pi, beams=beams, map="example_map")
    print("wrote", out, scans.shape)


# ----------------------------------------------------------------------------- rollouts
def _rollout(tag, map_name, num_agents, integrator_name, fov_name, n_envs, n_steps, poses0, actions,
             scan_every, seed=12345, params=None, timestep=0.01, ego_idx=0, agent_params=None, origin_yaw=0.0):
    """agent_params: {agent index: params dict} applied through F110Env.update_params(params, index) after
    construction (base_classes.py:507-527)."""
    gym = _ref()
    from f110_gym.envs.base_classes import Integrator
    integ = getattr(Integrator, integrator_name)
    fov = FOVS[fov_name]
    A = num_agents
    S = np.zeros((n_envs, n_steps + 1, A, 7))
    coll = np.zeros((n_envs, n_steps + 1, A))
    cidx = np.zeros((n_envs, n_steps + 1, A))
    tog = np.zeros((n_envs, n_steps + 1, A))
    lapc = np.zeros((n_envs, n_steps + 1, A))
    lapt = np.zeros((n_envs, n_steps + 1, A))
    done = np.zeros((n_envs, n_steps + 1), dtype=bool)
    obs_pose = np.zeros((n_envs, n_steps + 1, A, 5))   # poses_x, poses_y, poses_theta, linear_vels_x, ang_vels_z
    scan_steps = np.array(sorted(set([0, 1, 2] + list(range(0, n_steps + 1, scan_every)))))
    scans = np.zeros((n_envs, scan_steps.size, A, 1080))
    scan_sum = np.zeros((n_envs, n_steps + 1, A))
    scan_min = np.zeros((n_envs, n_steps + 1, A))
    kw = dict(map=rotated_map(map_name, origin_yaw) if origin_yaw else MAPS[map_name], map_ext=".png", num_agents=A,
              timestep=timestep, integrator=integ, fov=fov, seed=seed, ego_idx=ego_idx)
    if params is not None:
        kw["params"] = params
    for e in range(n_envs):
        env = gym.make("f110_gym:f110-v0", **kw)
        for a, p_a in (agent_params or {}).items():
            env.update_params(p_a, index=a)

        def rec(t, obs, d):
            for a in range(A):
                S[e, t, a] = env.sim.agents[a].state
                obs_pose[e, t, a] = [obs["poses_x"][a], obs["poses_y"][a], obs["poses_theta"][a],
                                     obs["linear_vels_x"][a], obs["ang_vels_z"][a]]
                scan_sum[e, t, a] = obs["scans"][a].sum()
                scan_min[e, t, a] = obs["scans"][a].min()
            coll[e, t] = obs["collisions"]
            cidx[e, t] = env.sim.collision_idx
            tog[e, t] = env.toggle_list
            lapc[e, t] = obs["lap_counts"]
            lapt[e, t] = obs["lap_times"]
            done[e, t] = d
            k = np.searchsorted(scan_steps, t)
            if k < scan_steps.size and scan_steps[k] == t:
                for a in range(A):
                    scans[e, k, a] = obs["scans"][a]

        obs, r, d, info = env.reset(poses0[e].copy())
        rec(0, obs, d)
        for t in range(n_steps):
            obs, r, d, info = env.step(actions[t, e].copy())
            rec(t + 1, obs, d)
    out = os.path.join(GOLDEN, tag + ".npz")
    extra = {}
    if origin_yaw:
        extra["origin_yaw"] = origin_yaw
    if agent_params:
        keys = sorted(next(iter(agent_params.values())).keys())
        extra.update(agent_param_idx=np.array(sorted(agent_params)), agent_param_keys=np.array(keys),
                     agent_param_vals=np.array([[agent_params[a][k] for k in keys] for a in sorted(agent_params)]))
    np.savez_compressed(out, map=map_name, num_agents=A, integrator=integrator_name, fov=fov, seed=seed,
                        timestep=timestep, ego_idx=ego_idx, **extra, poses0=poses0, actions=actions, states=S, obs_pose=obs_pose, collisions=coll,
                        collision_idx=cidx, toggles=tog, lap_counts=lapc, lap_times=lapt, done=done,
                        scan_steps=scan_steps, scans=scans, scan_sum=scan_sum, scan_min=scan_min)
    print("wrote", out, "collided envs:", int((coll.max(axis=(1, 2)) > 0).sum()), "of", n_envs,
          "max toggles", tog.max())


def _random_actions(n_steps, n_envs, A, seed=2024, speed_hi=8.0, hold=1):
    rng = np.random.default_rng(seed)
    n_draw = (n_steps + hold - 1) // hold
    steer = rng.uniform(-0.4189, 0.4189, (n_draw, n_envs, A))
    speed = rng.uniform(0.0, speed_hi, (n_draw, n_envs, A))
    act = np.stack([steer, speed], axis=-1)
    return np.repeat(act, hold, axis=0)[:n_steps]


def gen_rollout_a1_rk4():
    n, T = 8, 1000
    poses0 = waypoint_poses(n)[:, None, :]
    _rollout("rollout_a1_rk4", "example_map", 1, "RK4", "2pi", n, T, poses0, _random_actions(T, n, 1), 100)


def gen_rollout_a1_rk4_smooth():
    # piecewise-constant actions (held 25 steps): cars build speed, leave the kinematic regime,
    # some survive long enough to toggle the lap zone
    n, T = 6, 1000
    poses0 = waypoint_poses(n)[:, None, :]
    act = _random_actions(T, n, 1, seed=99, speed_hi=6.0, hold=25)
    act[..., 0] *= 0.35
    _rollout("rollout_a1_rk4_smooth", "example_map", 1, "RK4", "2pi", n, T, poses0, act, 250)


def gen_rollout_a1_euler():
    n, T = 4, 400
    rng = np.random.default_rng(5)
    poses0 = np.array([[0.0, 0.0, 1.2], [0.0, 0.0, -1.0], [0.3, -0.2, 4.5], [-0.2, 0.1, 0.2]])[:, None, :]
    act = _random_actions(T, n, 1, seed=11, speed_hi=5.0, hold=10)
    _rollout("rollout_a1_euler_berlin", "berlin", 1, "Euler", "4p7", n, T, poses0, act, 100, seed=777)


def gen_rollout_a2():
    # chaser 1.5 m behind the leader with a lateral offset; GJK + opponent ray-cast fire
    n, T = 8, 600
    rng = np.random.default_rng(4)
    lead = waypoint_poses(n)
    chase = lead.copy()
    lat = rng.uniform(-0.3, 0.3, n)
    chase[:, 0] += -1.5 * np.cos(lead[:, 2]) - lat * np.sin(lead[:, 2])
    chase[:, 1] += -1.5 * np.sin(lead[:, 2]) + lat * np.cos(lead[:, 2])
    poses0 = np.stack([lead, chase], axis=1)
    act = np.zeros((T, n, 2, 2))
    steer = rng.uniform(-0.15, 0.15, (T // 20 + 1, n, 2)).repeat(20, axis=0)[:T]
    act[..., 0] = steer
    act[:, :, 0, 1] = rng.uniform(0.0, 4.0, (T // 20 + 1, n)).repeat(20, axis=0)[:T]
    act[:, :, 1, 1] = rng.uniform(2.0, 8.0, (T // 20 + 1, n)).repeat(20, axis=0)[:T]
    _rollout("rollout_a2_rk4", "example_map", 2, "RK4", "2pi", n, T, poses0, act, 100)


def gen_rollout_a3():
    # three agents side by side on skirk: collision_multiple over 3 pairs (collision_models.py:185)
    n, T = 3, 300
    rng = np.random.default_rng(8)
    base = np.array([[0.0, 0.0, 1.5], [0.2, 0.1, 1.0], [-0.1, 0.0, 2.0]])
    poses0 = np.zeros((n, 3, 3))
    for e in range(n):
        th = base[e, 2]
        for a in range(3):
            off = (a - 1) * 0.45
            poses0[e, a] = [base[e, 0] - off * np.sin(th), base[e, 1] + off * np.cos(th) , th + 0.05 * a]
    act = np.zeros((T, n, 3, 2))
    act[..., 0] = rng.uniform(-0.2, 0.2, (T // 15 + 1, n, 3)).repeat(15, axis=0)[:T]
    act[..., 1] = rng.uniform(0.5, 4.0, (T // 15 + 1, n, 3)).repeat(15, axis=0)[:T]
    _rollout("rollout_a3_skirk", "skirk", 3, "RK4", "2pi", n, T, poses0, act, 100, seed=31)


def gen_rollout_a2_ego1():
    # the knobs the other rollouts leave at their defaults: ego_idx = 1 (done follows agent 1's collision flag and the
    # lap frame is agent 1's start pose, f110_env.py:292-299, 318-323), timestep 0.02, a heavier / shorter car for agent 0
    # set through update_params(index=0) (its own side_distances are NOT rebuilt by the reference: base_classes.py:116-156
    # runs once per process), vegas map
    n, T = 6, 400
    rng = np.random.default_rng(21)
    base = np.array([[0.0, 0.0, 0.0], [0.1, -0.1, 0.3], [-0.2, 0.1, -0.4], [0.0, 0.2, 3.0], [0.3, 0.0, 1.6], [-0.1, -0.2, 4.6]])
    poses0 = np.zeros((n, 2, 3))
    for e in range(n):
        th = base[e, 2]
        poses0[e, 0] = base[e]
        lat = rng.uniform(-0.25, 0.25)
        poses0[e, 1] = [base[e, 0] - 1.2 * np.cos(th) - lat * np.sin(th), base[e, 1] - 1.2 * np.sin(th) + lat * np.cos(th), th + 0.05]
    act = np.zeros((T, n, 2, 2))
    act[..., 0] = rng.uniform(-0.2, 0.2, (T // 10 + 1, n, 2)).repeat(10, axis=0)[:T]
    act[:, :, 0, 1] = rng.uniform(0.0, 3.0, (T // 10 + 1, n)).repeat(10, axis=0)[:T]
    act[:, :, 1, 1] = rng.uniform(1.0, 7.0, (T // 10 + 1, n)).repeat(10, axis=0)[:T]
    p0 = {'mu': 0.9, 'C_Sf': 4.2, 'C_Sr': 5.0, 'lf': 0.14, 'lr': 0.16, 'h': 0.08, 'm': 4.5, 'I': 0.05,
          's_min': -0.35, 's_max': 0.35, 'sv_min': -2.5, 'sv_max': 2.5, 'v_switch': 6.0, 'a_max': 7.0,
          'v_min': -4.0, 'v_max': 15.0, 'width': 0.28, 'length': 0.5}
    _rollout("rollout_a2_ego1_dt02_vegas", "vegas", 2, "RK4", "2pi", n, T, poses0, act, 100, seed=99,
             timestep=0.02, ego_idx=1, agent_params={0: p0})


def gen_rollout_a2_rot():
    # head-to-head pairs on example_map with the map origin rotated by -1.2 rad: the rotation of xy_2_rc under the whole
    # step (lidar, iTTC, opponent ray-cast, laps); same relative placement as rollout_a2_rk4
    n, T, yaw = 6, 400, -1.2
    rng = np.random.default_rng(40)
    lead = waypoint_poses(n)
    chase = lead.copy()
    lat = rng.uniform(-0.3, 0.3, n)
    chase[:, 0] += -1.5 * np.cos(lead[:, 2]) - lat * np.sin(lead[:, 2])
    chase[:, 1] += -1.5 * np.sin(lead[:, 2]) + lat * np.cos(lead[:, 2])
    poses0 = to_rotated_world(np.stack([lead, chase], axis=1), "example_map", yaw)
    act = np.zeros((T, n, 2, 2))
    act[..., 0] = rng.uniform(-0.15, 0.15, (T // 20 + 1, n, 2)).repeat(20, axis=0)[:T]
    act[:, :, 0, 1] = rng.uniform(0.0, 4.0, (T // 20 + 1, n)).repeat(20, axis=0)[:T]
    act[:, :, 1, 1] = rng.uniform(2.0, 8.0, (T // 20 + 1, n)).repeat(20, axis=0)[:T]
    _rollout("rollout_a2_rot_m1d2", "example_map", 2, "RK4", "2pi", n, T, poses0, act, 100, origin_yaw=yaw)


def gen_rollout_a1_rot():
    # single cars on berlin (resolution 0.05: the exact-division fallback of the cell index) with origin yaw 0.3, Euler
    n, T, yaw = 4, 300, 0.3
    base = np.array([[0.0, 0.0, 1.2], [0.0, 0.0, -1.0], [0.3, -0.2, 4.5], [-0.2, 0.1, 0.2]])
    poses0 = to_rotated_world(base, "berlin", yaw)[:, None, :]
    act = _random_actions(T, n, 1, seed=12, speed_hi=5.0, hold=10)
    _rollout("rollout_a1_rot_p0d3_berlin", "berlin", 1, "Euler", "4p7", n, T, poses0, act, 100, seed=778, origin_yaw=yaw)


def config2_stream(n_total=4096, n_steps=1000):
    """start poses k * R / 4096 on the raceline; per step per env steer ~ U(-0.4189, 0.4189),
    speed ~ U(0, 8) from default_rng(2024) (drawn as in tests/test_gpu_step.py::test_config2_*)."""
    poses = waypoint_poses(n_total)[:, None, :]
    rng = np.random.default_rng(2024)
    actions = np.stack([rng.uniform(-0.4189, 0.4189, (n_steps, n_total, 1)), rng.uniform(0.0, 8.0, (n_steps, n_total, 1))], axis=-1)
    return poses, actions


def gen_config2_slice():
    """256 of the 4096 envs of BASELINE config 2 (every 16th), all 1000 steps, through the unmodified reference.
    Compact: flags / toggles at every step, states every 50 steps, scan sums every 10, full scans of 32 envs at 3 steps."""
    gym = _ref()
    from f110_gym.envs.base_classes import Integrator
    stride, T = 16, 1000
    poses, actions = config2_stream(4096, T)
    poses, actions = poses[::stride], actions[:, ::stride]
    n = poses.shape[0]
    st_steps = np.arange(0, T + 1, 50)
    sum_steps = np.arange(0, T + 1, 10)
    sc_steps = np.array([0, 500, 1000])
    sc_envs = np.arange(0, n, 8)
    S = np.zeros((n, st_steps.size, 7))
    coll = np.zeros((n, T + 1), dtype=np.uint8)
    tog = np.zeros((n, T + 1), dtype=np.int8)
    done = np.zeros((n, T + 1), dtype=np.uint8)
    lapt = np.zeros((n, st_steps.size))
    ssum = np.zeros((n, sum_steps.size))
    scans = np.zeros((sc_envs.size, sc_steps.size, 1080))
    env = gym.make("f110_gym:f110-v0", map=MAPS["example_map"], map_ext=".png", num_agents=1, timestep=0.01,
                   integrator=Integrator.RK4, seed=12345)
    for e in range(n):
        def rec(t, obs, d):
            coll[e, t] = obs["collisions"][0]
            tog[e, t] = env.toggle_list[0]
            done[e, t] = d
            if t % 50 == 0:
                S[e, t // 50] = env.sim.agents[0].state
                lapt[e, t // 50] = obs["lap_times"][0]
            if t % 10 == 0:
                ssum[e, t // 10] = obs["scans"][0].sum()
            if e % 8 == 0 and t in (0, 500, 1000):
                scans[e // 8, (0, 500, 1000).index(t)] = obs["scans"][0]
        obs, r, d, info = env.reset(poses[e].copy())        # re-seeds the noise stream, clears every counter (f110_env.py:304-347)
        rec(0, obs, d)
        for t in range(T):
            obs, r, d, info = env.step(actions[t, e].copy())
            rec(t + 1, obs, d)
    out = os.path.join(GOLDEN, "config2_slice.npz")
    np.savez_compressed(out, stride=stride, n_total=4096, state_steps=st_steps, states=S, lap_times=lapt, collisions=coll,
                        toggles=tog, done=done, sum_steps=sum_steps, scan_sum=ssum, scan_steps=sc_steps, scan_envs=sc_envs, scans=scans)
    print("wrote", out, "envs that hit a wall:", int((coll.max(axis=1) > 0).sum()), "of", n)


def gen_rollout_a2_fov():
    # two agents with a 4.7 rad field of view and the Euler integrator (the other A >= 2 goldens are 2 pi / RK4): the opponent's
    # corners can lie outside the field of view, where get_blocked_view_indices (laser_models.py:283-315) clamps to the first /
    # last beam; chaser behind, beside and in front of the leader
    n, T = 6, 300
    rng = np.random.default_rng(52)
    base = np.array([[0.0, 0.0, 1.5], [0.2, 0.1, 1.0], [-0.1, 0.0, 2.0], [0.0, 0.3, -1.2], [0.3, -0.3, 0.4], [0.1, 0.1, 3.0]])
    rel = np.array([[-1.2, 0.1], [-0.2, 0.5], [1.0, -0.2], [-0.9, -0.4], [0.1, -0.55], [-1.5, 0.0]])      # chaser in the leader's frame
    poses0 = np.zeros((n, 2, 3))
    for e in range(n):
        th = base[e, 2]
        poses0[e, 0] = base[e]
        poses0[e, 1] = [base[e, 0] + rel[e, 0] * np.cos(th) - rel[e, 1] * np.sin(th), base[e, 1] + rel[e, 0] * np.sin(th) + rel[e, 1] * np.cos(th),
                        th + rng.uniform(-0.2, 0.2)]
    act = np.zeros((T, n, 2, 2))
    act[..., 0] = rng.uniform(-0.25, 0.25, (T // 12 + 1, n, 2)).repeat(12, axis=0)[:T]
    act[:, :, 0, 1] = rng.uniform(0.0, 3.0, (T // 12 + 1, n)).repeat(12, axis=0)[:T]
    act[:, :, 1, 1] = rng.uniform(1.0, 6.0, (T // 12 + 1, n)).repeat(12, axis=0)[:T]
    _rollout("rollout_a2_euler_fov4p7_skirk", "skirk", 2, "Euler", "4p7", n, T, poses0, act, 100, seed=41)


def gen_waypoint_follow():
    """BASELINE config 1: examples/waypoint_follow.py with render() skipped (GL unavailable on this box)."""
    gym = _ref()
    sys.path.insert(0, REF + "/examples")
    import yaml
    from argparse import Namespace
    import waypoint_follow as wf
    from f110_gym.envs.base_classes import Integrator
    os.chdir(REF + "/examples")
    with open("config_example_map.yaml") as f:
        conf = Namespace(**yaml.load(f, Loader=yaml.FullLoader))
    planner = wf.PurePursuitPlanner(conf, 0.17145 + 0.15875)
    env = gym.make("f110_gym:f110-v0", map=conf.map_path, map_ext=conf.map_ext, num_agents=1, timestep=0.01,
                   integrator=Integrator.RK4)
    obs, r, done, info = env.reset(np.array([[conf.sx, conf.sy, conf.stheta]]))
    acts, poses, vels, laps, lapt, tog = [], [], [], [], [], []
    poses.append([obs["poses_x"][0], obs["poses_y"][0], obs["poses_theta"][0]])
    while not done:
        speed, steer = planner.plan(obs["poses_x"][0], obs["poses_y"][0], obs["poses_theta"][0],
                                    0.82461887897713965, 1.375)
        obs, r, done, info = env.step(np.array([[steer, speed]]))
        acts.append([steer, speed])
        poses.append([obs["poses_x"][0], obs["poses_y"][0], obs["poses_theta"][0]])
        vels.append(obs["linear_vels_x"][0])
        laps.append(obs["lap_counts"][0]); lapt.append(obs["lap_times"][0]); tog.append(env.toggle_list[0])
    out = os.path.join(GOLDEN, "waypoint_follow.npz")
    np.savez_compressed(out, actions=np.array(acts), poses=np.array(poses), vels=np.array(vels),
                        lap_counts=np.array(laps), lap_times=np.array(lapt), toggles=np.array(tog),
                        final_collision=obs["collisions"][0], final_scan=obs["scans"][0],
                        start=np.array([conf.sx, conf.sy, conf.stheta]))
    print("wrote", out, "steps", len(acts), "lap time", lapt[-1], "collision", obs["collisions"][0])


def gen_kats():
    """Known-answer inputs of the reference's embedded unit tests, evaluated by the reference itself."""
    _ref()
    from f110_gym.envs import collision_models as cm, dynamic_models as dm, laser_models as lm
    out = {}
    # CollisionTests.setUp / test_multiple_collisions (collision_models.py:272-324)
    np.random.seed(1234)
    v1 = np.asarray([[4, 11.], [5, 5], [9, 9], [10, 10]])
    bodies = [v1 + np.random.normal(size=v1.shape) / 100. for _ in range(6)] + [v1 + 10.]
    allv = np.stack(bodies)
    c, ci = cm.collision_multiple(allv)
    out.update(gjk_bodies=allv, gjk_collisions=c, gjk_collision_idx=ci)
    # 1000 jittered pairs (test_random_collision :306-311) + far/near separated pairs
    rng = np.random.default_rng(3)
    pa = v1 + rng.normal(size=(1000, 4, 2)) / 100.
    pb = v1 + rng.normal(size=(1000, 4, 2)) / 100.
    pb[500:] += rng.uniform(-8, 8, size=(500, 1, 2))
    out.update(gjk_pairs_a=pa, gjk_pairs_b=pb,
               gjk_pairs_hit=np.array([cm.collision(np.ascontiguousarray(pa[i]), np.ascontiguousarray(pb[i])) for i in range(1000)]))
    # car rectangles at random relative poses (get_vertices :238 + collision :114)
    poses_a = np.concatenate([rng.uniform(-1, 1, (2000, 2)), rng.uniform(-4, 4, (2000, 1))], axis=1)
    poses_b = np.concatenate([rng.uniform(-1, 1, (2000, 2)), rng.uniform(-4, 4, (2000, 1))], axis=1)
    poses_b[:50] = poses_a[:50]                       # identical centres: d==0 branch (:133-134)
    poses_b[50:100, 2] = poses_a[50:100, 2]           # parallel
    va = np.stack([cm.get_vertices(p, 0.58, 0.31) for p in poses_a])
    vb = np.stack([cm.get_vertices(p, 0.58, 0.31) for p in poses_b])
    hit = np.array([cm.collision(np.ascontiguousarray(va[i]), np.ascontiguousarray(vb[i])) for i in range(2000)])
    out.update(car_poses_a=poses_a, car_poses_b=poses_b, car_verts_a=va, car_verts_b=vb, car_hit=hit)
    # opponent ray-cast (laser_models.py:319) on a flat 30 m scan
    fov = 2 * np.pi
    angles = np.array([-fov / 2. + i * (fov / 1079) for i in range(1080)])
    rc_scans = []
    ego = np.array([0.3, -0.2, 0.7])
    for i in range(64):
        scan = np.full(1080, 30.0)
        rc_scans.append(lm.ray_cast(ego, scan, angles, va[i] + np.array([ego[0], ego[1]])))
    out.update(rc_ego=ego, rc_scans=np.stack(rc_scans), rc_angles=angles)
    # check_ttc_jit (laser_models.py:189)
    # dynamics known answers: evaluate the RHS on random states (both regimes) and pid
    p = [1.0489, 4.718, 5.4562, 0.15875, 0.17145, 0.074, 3.74, 0.04712, -0.4189, 0.4189, -3.2, 3.2, 7.319, 9.51, -5.0, 20.0]
    X = rng.uniform(-1, 1, (512, 7)) * np.array([10, 10, 0.45, 12, 6.3, 3, 0.5])
    X[:64, 3] = rng.uniform(-0.6, 0.6, 64)
    U = rng.uniform(-1, 1, (512, 2)) * np.array([4.0, 12.0])
    F = np.stack([dm.vehicle_dynamics_st(X[i], U[i], *p) for i in range(512)])
    pid_in = rng.uniform(-1, 1, (512, 4)) * np.array([10, 0.5, 10, 0.5])
    pid_out = np.array([dm.pid(s[0], s[1], s[2], s[3], 3.2, 9.51, 20.0, -5.0) for s in pid_in])
    out.update(dyn_params=np.array(p), dyn_x=X, dyn_u=U, dyn_f=F, pid_in=pid_in, pid_out=pid_out)
    path = os.path.join(GOLDEN, "kats.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, "car hits", int(hit.sum()))


SCENARIOS = {
    "kats": gen_kats,
    "rollout_a1_rk4": gen_rollout_a1_rk4,
    "rollout_a1_rk4_smooth": gen_rollout_a1_rk4_smooth,
    "rollout_a1_euler": gen_rollout_a1_euler,
    "rollout_a2": gen_rollout_a2,
    "rollout_a3": gen_rollout_a3,
    "rollout_a2_ego1": gen_rollout_a2_ego1,
    "rollout_a2_rot": gen_rollout_a2_rot,
    "rollout_a2_fov": gen_rollout_a2_fov,
    "rollout_a1_rot": gen_rollout_a1_rot,
    "config2_slice": gen_config2_slice,
    "waypoint_follow": gen_waypoint_follow,
}
for _m in MAPS:
    for _f in FOVS:
        SCENARIOS["scans_%s_%s" % (_m, _f)] = (lambda m=_m, f=_f: gen_scans(m, f))
SCENARIOS["scans_example_map_2pi_b270"] = lambda: gen_scans("example_map", "2pi", beams=270, n_poses=12)
SCENARIOS["scans_example_map_2pi_b2160"] = lambda: gen_scans("example_map", "2pi", beams=2160, n_poses=6)
SCENARIOS["scans_example_map_2pi_b20001"] = gen_scans_many_beams
# rotated map origins (the ROT instantiation of the ray-cast had no reference vector)
SCENARIOS["scans_berlin_2pi_rot_p0d3"] = lambda: gen_scans("berlin", "2pi", yaw=0.3)
SCENARIOS["scans_berlin_4p7_rot_m1d2"] = lambda: gen_scans("berlin", "4p7", yaw=-1.2)
SCENARIOS["scans_example_map_2pi_rot_p0d3"] = lambda: gen_scans("example_map", "2pi", yaw=0.3)
SCENARIOS["scans_example_map_2pi_rot_m1d2"] = lambda: gen_scans("example_map", "2pi", yaw=-1.2)
SCENARIOS["scans_example_map_2pi_b2160_rot_m1d2"] = lambda: gen_scans("example_map", "2pi", beams=2160, n_poses=6, yaw=-1.2)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", nargs="*", default=None)
    ap.add_argument("--worker", default=None)
    args = ap.parse_args()
    os.makedirs(GOLDEN, exist_ok=True)
    if args.worker:
        SCENARIOS[args.worker]()
        return
    names = args.only or list(SCENARIOS)
    for n in names:
        subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", n], check=True)


if __name__ == "__main__":
    main()
