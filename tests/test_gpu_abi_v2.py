"""ABI v2 additions on the GPU: the compact uint16 scan format, Simulator.reset without a step, the constraint functions,
a failed map update leaving the handle intact, non-default ScanSimulator2D arguments, the L2 probe."""
import ctypes
import os

import numpy as np
import pytest

import f110_oracle as fo
from conftest import GOLDEN, MAPS

pytestmark = pytest.mark.gpu


def _actions(torch, N, A, g):
    act = torch.rand((N, A, 2), generator=g, device="cuda", dtype=torch.float64)
    act[..., 0] = (act[..., 0] * 2 - 1) * 0.4189
    act[..., 1] = act[..., 1] * 8.0
    return act


@pytest.mark.parametrize("N", [700, 6000, 40000])
def test_u16_scan_format_is_the_float_scan_rounded_to_half_millimetres(N):
    """F110_SCAN_U16 (fp32 path, A = 1): the scan is stored as round(2000 r) in uint16 -- |error| <= 0.25 mm, a noisy range
    below zero becomes 0 -- while iTTC sees the unrounded range: states, flags and resets are bit-identical to a native
    handle.  Sizes: CTA-per-scan kernel, warp-per-scan kernel, spatially ordered launch."""
    import torch
    from red_gym_b200 import BatchF110Env
    from red_gym_b200.host import raceline_poses
    pool = raceline_poses(256)[:, None, :]
    kw = dict(num_envs=N, num_agents=1, map=os.path.join(MAPS, "example_map"), dtype="fp32", noise="philox", auto_reset=True, reset_poses=pool)
    a, b = BatchF110Env(**kw), BatchF110Env(scan_format="u16", **kw)
    assert b.scans.dtype == torch.uint16 and tuple(b.scans.shape) == (N, 1, 1080)
    poses = raceline_poses(N)[:, None, :]
    a.reset(poses); b.reset(poses)
    g = torch.Generator(device="cuda").manual_seed(3)
    for t in range(120):
        act = _actions(torch, N, 1, g)
        a.step(act); b.step(act)
    assert torch.equal(a.obs_block, b.obs_block) and torch.equal(a.done, b.done) and torch.equal(a.episode_count, b.episode_count)
    assert int(a.episode_count.max()) >= 2
    want = torch.clamp(a.scans.double() * 2000.0, min=0.0)
    got = b.scans.to(torch.int32).double()
    assert float((got - want).abs().max()) <= 0.5 + 4e-3            # round to nearest LSB; the product r * 2000 is fp32 (60000 * 2^-24 = 0.0036)
    assert float((got / 2000.0 - a.scans.double().clamp(min=0)).abs().max()) <= 0.25e-3 + 2e-6
    # through f110_step_host: the compact scans come back as 2 bytes per beam
    act_h = torch.zeros((N, 1, 2), dtype=torch.float64).pin_memory()
    obs_h = [torch.zeros((10, N, 1), dtype=torch.float64).pin_memory() for _ in range(2)]
    sc_a = torch.zeros((N, 1, 1080), dtype=torch.float32).pin_memory()
    sc_b = torch.zeros((N, 1, 1080), dtype=torch.uint16).pin_memory()
    done_h = [torch.zeros((N,), dtype=torch.uint8).pin_memory() for _ in range(2)]
    act_h.copy_(_actions(torch, N, 1, g))
    a.step_host(act_h, obs_h[0], sc_a, done_h[0]); b.step_host(act_h, obs_h[1], sc_b, done_h[1])
    assert torch.equal(obs_h[0], obs_h[1]) and torch.equal(done_h[0], done_h[1])
    assert torch.equal(sc_b.cuda(), b.scans) and torch.equal(sc_a.cuda(), a.scans)
    assert float((sc_b.to(torch.int32).double() / 2000.0 - sc_a.double().clamp(min=0)).abs().max()) <= 0.25e-3 + 2e-6
    # f110_raycast of a u16 handle still returns float ranges
    assert torch.equal(a.raycast(poses[:50, 0]), b.raycast(poses[:50, 0]))


def test_u16_needs_the_fp32_single_agent_path():
    from red_gym_b200 import BatchF110Env
    kw = dict(num_envs=4, map=os.path.join(MAPS, "example_map"), scan_format="u16")
    with pytest.raises(ValueError):
        BatchF110Env(num_agents=1, dtype="fp64", **kw)
    with pytest.raises(ValueError):
        BatchF110Env(num_agents=2, dtype="fp32", **kw)
    with pytest.raises(ValueError):
        BatchF110Env(num_agents=1, dtype="fp32", num_beams=271, **kw)


def test_simulator_reset_takes_no_step():
    """Simulator.reset (base_classes.py:607-623) only resets the agents and returns None; the zero-action step is
    F110Env.reset's (f110_env.py:335-336).  Driving the Simulator directly therefore reproduces the reference's rollout
    with reset + step(zeros) + steps, and obs['ego_idx'] is the Simulator's own (0 under F110Env, whatever the env's)."""
    import f110_gym
    from f110_gym.envs.base_classes import Integrator, RaceCar, Simulator
    from red_gym_b200.host import DEFAULT_PARAMS
    d = np.load(os.path.join(GOLDEN, "rollout_a2_rk4.npz"))
    sim = Simulator(dict(DEFAULT_PARAMS), 2, 12345, 2 * np.pi, time_step=0.01, integrator=Integrator.RK4)
    with pytest.raises(ValueError, match="Map is not set"):
        sim.step(np.zeros((2, 2)))
    sim.set_map(os.path.join(MAPS, "example_map.yaml"), ".png")
    assert sim.reset(d["poses0"][0]) is None
    assert isinstance(sim.agents[0], RaceCar) and len(sim.agents) == 2 and sim.agents[0].is_ego and not sim.agents[1].is_ego
    st = sim.agents[1].state
    assert st.shape == (7,) and st[0] == d["poses0"][0, 1, 0] and st[4] == d["poses0"][0, 1, 2] and (st[2:4] == 0).all()
    assert sim.agents[1].in_collision is False and sim.agents[0].scan_angles.shape == (1080,)
    obs = sim.step(np.zeros((2, 2)))                       # what F110Env.reset does next
    for a in range(2):
        assert abs(sim.agents[a].state - d["states"][0, 0, a]).max() <= 1e-9
        assert np.abs(obs["scans"][a] - d["scans"][0, 0, a]).max() <= 1e-4
    for t in range(100):
        obs = sim.step(d["actions"][t, 0])
        assert (obs["collisions"] == d["collisions"][0, t + 1]).all()
    assert abs(sim.agents[0].state - d["states"][0, 100, 0]).max() <= 1e-9 * 100
    with pytest.raises(ValueError, match="Number of poses"):
        sim.reset(np.zeros((3, 3)))
    # ego_idx quirk: env with ego_idx = 1 -> done / laps follow agent 1, the observation still says 0 (f110_env.py:190)
    e = np.load(os.path.join(GOLDEN, "rollout_a2_ego1_dt02_vegas.npz"))
    env = f110_gym.make('f110_gym:f110-v0', map=os.path.join(MAPS, "vegas"), num_agents=2, timestep=0.02, ego_idx=1, seed=99)
    p0 = dict(zip([str(k) for k in e["agent_param_keys"]], e["agent_param_vals"][0]))
    env.update_params(p0, index=0)
    obs, r, done, info = env.reset(e["poses0"][1])
    assert obs["ego_idx"] == 0 and r == 0.02
    for t in range(150):
        obs, r, done, info = env.step(e["actions"][t, 1])
        assert done == e["done"][1, t + 1] and (obs["collisions"] == e["collisions"][1, t + 1]).all()
        assert (env.toggle_list == e["toggles"][1, t + 1]).all()
    assert abs(obs["poses_x"][1] - e["obs_pose"][1, 150, 1, 0]) <= 1e-9 * 10


def test_constraint_functions():
    """accl_constraints / steering_constraint (dynamic_models.py:30-87): the device functions of the step behind the
    drop-in module, against the reference's branch table."""
    from f110_gym.envs.dynamic_models import accl_constraints, steering_constraint
    from red_gym_b200 import models
    v_switch, a_max, v_min, v_max = 7.319, 9.51, -5.0, 20.0
    assert accl_constraints(1.0, 3.0, v_switch, a_max, v_min, v_max) == 3.0
    assert accl_constraints(1.0, 30.0, v_switch, a_max, v_min, v_max) == a_max
    assert accl_constraints(1.0, -30.0, v_switch, a_max, v_min, v_max) == -a_max
    assert accl_constraints(10.0, 30.0, v_switch, a_max, v_min, v_max) == a_max * v_switch / 10.0
    assert accl_constraints(-5.0, -1.0, v_switch, a_max, v_min, v_max) == 0.0 and accl_constraints(20.0, 1.0, v_switch, a_max, v_min, v_max) == 0.0
    assert steering_constraint(0.1, 1.0, -0.4189, 0.4189, -3.2, 3.2) == 1.0
    assert steering_constraint(0.1, 9.0, -0.4189, 0.4189, -3.2, 3.2) == 3.2 and steering_constraint(0.1, -9.0, -0.4189, 0.4189, -3.2, 3.2) == -3.2
    assert steering_constraint(0.4189, 1.0, -0.4189, 0.4189, -3.2, 3.2) == 0.0 and steering_constraint(-0.5, -1.0, -0.4189, 0.4189, -3.2, 3.2) == 0.0
    rng = np.random.default_rng(0)
    vel, acc = rng.uniform(-7, 25, 4000), rng.uniform(-15, 15, 4000)
    got = models.accl_constraints(vel, acc, v_switch, a_max, v_min, v_max).cpu().numpy()
    pos = np.where(vel > v_switch, a_max * v_switch / np.where(vel > v_switch, vel, 1.0), a_max)
    want = np.where(((vel <= v_min) & (acc <= 0)) | ((vel >= v_max) & (acc >= 0)), 0.0, np.where(acc <= -a_max, -a_max, np.where(acc >= pos, pos, acc)))
    assert (got == want).all()


def test_failed_map_update_keeps_the_previous_map():
    """set_map is transactional: a rejected update (here: the padded fp32 layout of a huge fine-grained map would exceed
    2^31 cells) must leave the handle stepping on the map it had, not on freed memory."""
    from red_gym_b200 import BatchF110Env, host
    env = BatchF110Env(num_envs=64, num_agents=1, map=os.path.join(MAPS, "berlin"), dtype="fp32", noise="off")
    poses = np.zeros((64, 1, 3))
    env.reset(poses)
    before = env.scans.clone()
    bad = host.MapData.from_free_mask(np.ones((3000, 3000), dtype=np.uint8), 0.0005)      # pad = 60002 cells per side
    with pytest.raises(ValueError, match="too large"):
        env._upload_map(bad)
    env.reset(poses)
    assert (env.scans == before).all()
    with pytest.raises(ValueError, match="eps"):
        BatchF110Env(num_envs=1, num_agents=1, map=os.path.join(MAPS, "berlin"), dtype="fp32", noise="off", eps=0.06)


def test_scan_simulator_takes_non_default_arguments():
    """ScanSimulator2D(num_beams, fov, eps, theta_dis, max_range) (laser_models.py:360): values other than the defaults."""
    from f110_gym.envs.laser_models import ScanSimulator2D
    sim = ScanSimulator2D(540, 4.0, eps=0.03, theta_dis=1500, max_range=12.0)
    sim.set_map(os.path.join(MAPS, "berlin.yaml"), ".png")
    poses = np.array([[0.0, 0.0, 0.3], [0.5, -0.2, 2.0], [-0.3, 0.1, -1.0]])
    got = np.stack([sim.scan(p, None) for p in poses])
    L = fo.lib()
    h = L.f110o_create(1, 1, 540, 1, 0, 1500, 0.01, 4.0, 0.03, 12.0, 0.005)
    m = fo.load_map(os.path.join(MAPS, "berlin.yaml"), ".png")
    L.f110o_set_map(h, fo._p(m["dt"]), m["h"], m["w"], m["res"], m["ox"], m["oy"], m["oc"], m["os"])
    th = np.linspace(0.0, 2 * np.pi, num=1500)
    s, c = np.sin(th), np.cos(th)
    sa, bc, sd = fo.beam_tables(540, 4.0, fo.DEFAULT_PARAMS)
    L.f110o_set_tables(h, fo._p(s), fo._p(c), fo._p(sa), fo._p(bc), fo._p(sd))
    want = np.empty((3, 540)); lk = np.zeros(3, dtype=np.uint64)
    L.f110o_scan(h, fo._p(np.ascontiguousarray(poses)), 3, fo._p(want), fo._p(lk), 0)
    assert (got == want).all() and got.max() <= 12.0
    L.f110o_destroy(h)


def test_observation_views_alias_and_copy():
    import torch
    from red_gym_b200 import BatchF110Env
    from red_gym_b200.host import raceline_poses
    N = 32
    env = BatchF110Env(num_envs=N, num_agents=1, map=os.path.join(MAPS, "example_map"))
    obs0, _, _, _ = env.reset(raceline_poses(N)[:, None, :], copy=True)
    act = torch.zeros((N, 1, 2), dtype=torch.float64, device="cuda")
    act[..., 1] = 3.0
    view, _, _, _ = env.step(act)
    kept = view["poses_x"].clone()
    for _ in range(5):
        env.step(act)
    assert not torch.equal(view["poses_x"], kept)             # a view follows the simulator ...
    assert torch.equal(obs0["poses_x"].cpu(), torch.from_numpy(raceline_poses(N)[:, None, 0]))      # ... a copy does not
    env.enable_timing(True)
    h = [torch.zeros((N, 1, 2), dtype=torch.float64).pin_memory(), torch.zeros((10, N, 1), dtype=torch.float64).pin_memory()]
    env.step_host(h[0], h[1])
    assert env.last_raycast_ms() > 0.0
    env.enable_timing(False)
    with pytest.raises(ValueError, match="no timed"):
        env.last_raycast_ms()
    env.close()
    assert not hasattr(env, "scans")


def test_l2_probe_reports_a_bandwidth():
    from red_gym_b200 import _native as nat
    out = (ctypes.c_double * 6)()
    nat.check(nat.load().f110_l2_probe(0, 26 << 20, 3, out, None))
    stream, gather = out[0], out[1]
    assert 1000.0 < stream < 60000.0                       # GB/s
    assert 1e10 < gather < 2e12                            # 32-byte sectors per second


def test_noise_rows_grown_lazily_match_rows_reserved_up_front():
    """Parity mode: step() extends the noise table 128 rows at a time, uploaded beside the stepping stream
    (f110_set_noise_rows: rows behind the table's end do not drain the caller's stream); the scans must be the ones of an env
    whose rows were all uploaded before the first step -- across five chunk borders and two table reallocations, on a
    stream of the caller's own with steps queued up behind each other."""
    import torch
    from red_gym_b200 import BatchF110Env
    from red_gym_b200.host import raceline_poses
    N, T = 600, 700
    kw = dict(num_envs=N, num_agents=1, map=os.path.join(MAPS, "example_map"), dtype="fp64", noise="numpy", seed=4242)
    poses = raceline_poses(N)[:, None, :]
    g = torch.Generator(device="cuda").manual_seed(11)
    acts = [_actions(torch, N, 1, g) for _ in range(16)]
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        a, b = BatchF110Env(**kw), BatchF110Env(**kw)
        a.reset(poses); b.reset(poses)
        b.reserve_noise(T + 2)
        assert b._noise_rows >= T + 2 and a._noise_rows < 300
        for t in range(T):
            a.step(acts[t % 16]); b.step(acts[t % 16])
            if t % 97 == 0 or t == T - 1:
                assert torch.equal(a.scans, b.scans) and torch.equal(a.obs_block, b.obs_block), t
        assert T <= a._noise_rows < T + 130       # grown in chunks of 128 rows, not doubled
        assert torch.equal(a.noise_step, b.noise_step)
    torch.cuda.synchronize()
