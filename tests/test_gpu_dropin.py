"""The drop-in `f110_gym` package: same API, types and results as the reference's F110Env."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, MAPS

pytestmark = pytest.mark.gpu


def test_waypoint_follow_lap_reproduces_reference():
    """BASELINE config 1 (examples/waypoint_follow.py): replay the recorded pure-pursuit actions."""
    import f110_gym
    from f110_gym.envs.base_classes import Integrator
    d = np.load(os.path.join(GOLDEN, "waypoint_follow.npz"))
    env = f110_gym.make('f110_gym:f110-v0', map=os.path.join(MAPS, "example_map"), map_ext='.png', num_agents=1,
                        timestep=0.01, integrator=Integrator.RK4)
    obs, step_reward, done, info = env.reset(np.array([d["start"]]))
    assert set(obs.keys()) == {'ego_idx', 'scans', 'poses_x', 'poses_y', 'poses_theta', 'linear_vels_x', 'linear_vels_y',
                               'ang_vels_z', 'collisions', 'lap_times', 'lap_counts'}
    assert isinstance(obs['poses_x'], list) and isinstance(obs['poses_x'][0], np.float64)
    assert isinstance(obs['scans'], list) and obs['scans'][0].shape == (1080,) and obs['scans'][0].dtype == np.float64
    assert obs['collisions'].shape == (1,) and obs['lap_times'].shape == (1,) and step_reward == 0.01 and done is False
    T = d["actions"].shape[0]
    laptime = 0.0
    for t in range(T):
        obs, step_reward, done, info = env.step(d["actions"][t].reshape(1, 2))
        laptime += step_reward
        p = d["poses"][t + 1]
        assert abs(obs['poses_x'][0] - p[0]) <= 1e-9 * max(1, abs(p[0])) and abs(obs['poses_y'][0] - p[1]) <= 1e-9 * max(1, abs(p[1]))
        assert obs['lap_counts'][0] == d["lap_counts"][t]
        if t < T - 1:
            assert not done
    assert done and obs['collisions'][0] == 0.0 and obs['lap_counts'][0] == 2.0
    assert abs(obs['lap_times'][0] - d["lap_times"][-1]) < 1e-9
    assert np.abs(obs['scans'][0] - d["final_scan"]).max() <= 1e-4
    assert info['checkpoint_done'].all()
    env.render('human_fast')
    with pytest.raises(AssertionError):
        env.render('rgb')


def test_two_agent_env_and_errors():
    import f110_gym
    from f110_gym.envs.base_classes import Integrator
    d = np.load(os.path.join(GOLDEN, "rollout_a2_rk4.npz"))
    env = f110_gym.make('f110_gym:f110-v0', map=os.path.join(MAPS, "example_map"), num_agents=2)   # default integrator RK4, fov 2pi
    obs, r, done, info = env.reset(d["poses0"][0])
    for t in range(120):
        obs, r, done, info = env.step(d["actions"][t, 0])
        assert (obs['collisions'] == d["collisions"][0, t + 1]).all()
        assert (env.sim.collision_idx == d["collision_idx"][0, t + 1]).all()
        assert done == d["done"][0, t + 1]
    assert abs(obs['poses_x'][1] - d["obs_pose"][0, 120, 1, 0]) < 1e-9
    with pytest.raises(ValueError, match="Number of poses"):
        env.reset(np.zeros((3, 3)))
    with pytest.raises(IndexError):
        env.update_params(env.params, index=2)
    with pytest.raises(SyntaxError):
        f110_gym.make('f110_gym:f110-v0', map=os.path.join(MAPS, "example_map"), num_agents=1, integrator="Heun")
    assert Integrator.RK4.value == 1 and Integrator.Euler.value == 2
