"""Drop-in F110Env (reference: gym/f110_gym/envs/f110_env.py:53-416) backed by the B200 library.

Same kwargs, same `reset(poses) / step(action) / render(mode)` signatures, same observation dict
(python lists of np.float64, list of (1080,) float64 scans, (A,) float64 arrays) and the same
exceptions.  One env is a `num_envs=1` view of the batched simulator; use
`red_gym_b200.BatchF110Env` to step many envs at once.
"""
import os
import time

import numpy as np

try:   # the reference derives from gym.Env; keep that when gym exists
    import gym
    _Base = gym.Env
except Exception:
    _Base = object

from f110_gym.envs.base_classes import Simulator, Integrator
from red_gym_b200 import host

# rendering constants of the reference (f110_env.py:48-51)
VIDEO_W = 600
VIDEO_H = 400
WINDOW_W = 1000
WINDOW_H = 800


class F110Env(_Base):
    metadata = {'render.modes': ['human', 'human_fast']}

    # class-level like the reference (f110_env.py:96-98)
    renderer = None
    current_obs = None
    render_callbacks = []

    def __init__(self, **kwargs):
        self.seed = kwargs.get('seed', 12345)
        self.map_name = kwargs.get('map', None)
        self.map_path = host.resolve_map(self.map_name)
        self.map_ext = kwargs.get('map_ext', '.png')
        self.params = kwargs.get('params', dict(host.DEFAULT_PARAMS))
        self.num_agents = kwargs.get('num_agents', 2)
        self.timestep = kwargs.get('timestep', 0.01)
        self.ego_idx = kwargs.get('ego_idx', 0)
        self.integrator = kwargs.get('integrator', Integrator.RK4)
        self.sim_car_fov = kwargs.get('fov', 2 * np.pi)
        self.start_thresh = 0.5

        self.poses_x = []
        self.poses_y = []
        self.poses_theta = []
        self.collisions = np.zeros((self.num_agents, ))
        self.lap_times = np.zeros((self.num_agents, ))
        self.lap_counts = np.zeros((self.num_agents, ))
        self.current_time = 0.0
        self._toggles = np.zeros((self.num_agents,))
        self.start_xs = np.zeros((self.num_agents, ))
        self.start_ys = np.zeros((self.num_agents, ))
        self.start_thetas = np.zeros((self.num_agents, ))
        self.start_rot = np.eye(2)

        # like the reference (f110_env.py:190) the Simulator is not told the env's ego_idx: obs['ego_idx'] stays 0;
        # the env-level ego (done flag, lap frame) goes to the library separately
        self.sim = Simulator(self.params, self.num_agents, self.seed, self.sim_car_fov, time_step=self.timestep,
                             integrator=self.integrator, device=kwargs.get('device', 0), env_ego_idx=self.ego_idx)
        self.sim.set_map(self.map_path, self.map_ext)
        self.render_obs = None

    def _finish(self, obs):
        # the device already did the time/lap bookkeeping of f110_env.py:292-299 (_check_done :202-244)
        self.lap_times = self.sim.lap_times
        self.lap_counts = self.sim.lap_counts
        obs['lap_times'] = self.lap_times
        obs['lap_counts'] = self.lap_counts
        F110Env.current_obs = obs
        self.render_obs = {k: obs[k] for k in ('ego_idx', 'poses_x', 'poses_y', 'poses_theta', 'lap_times', 'lap_counts')}
        self.poses_x = obs['poses_x']
        self.poses_y = obs['poses_y']
        self.poses_theta = obs['poses_theta']
        self.collisions = obs['collisions']
        self._toggles = None                  # read back from the device only if somebody looks (toggle_list below)
        done = self.sim.done
        info = {'checkpoint_done': self.lap_counts >= 2}      # toggle_list >= 4 (f110_env.py:299); lap_counts = toggles // 2
        return obs, self.timestep, bool(done), info

    @property
    def toggle_list(self):
        """Start-line toggles per agent (f110_env.py:183, 233-240), kept on the device; fetched on access."""
        if self._toggles is None:
            self._toggles = self.sim._batch.toggle_list.cpu().numpy()[0].astype(np.float64)
        return self._toggles

    @toggle_list.setter
    def toggle_list(self, value):
        self._toggles = np.asarray(value, dtype=np.float64)

    def step(self, action):
        obs = self.sim.step(action)
        self.current_time = self.current_time + self.timestep
        return self._finish(obs)

    def reset(self, poses):
        poses = np.asarray(poses, dtype=np.float64)
        self.current_time = 0.0
        self.start_xs = poses[:, 0]
        self.start_ys = poses[:, 1]
        self.start_thetas = poses[:, 2]
        th = self.start_thetas[self.ego_idx]
        self.start_rot = np.array([[np.cos(-th), -np.sin(-th)], [np.sin(-th), np.cos(-th)]])
        self.sim.reset(poses)
        # get no input observations (f110_env.py:335-336)
        return self.step(np.zeros((self.num_agents, 2)))

    def update_map(self, map_path, map_ext):
        self.sim.set_map(map_path, map_ext)

    def update_params(self, params, index=-1):
        self.sim.update_params(params, agent_idx=index)

    def add_render_callback(self, callback_func):
        F110Env.render_callbacks.append(callback_func)

    def render(self, mode='human'):
        """The reference draws with pyglet/OpenGL (f110_env.py:385-416) -- unavailable on this box (no GL/X).
        'human' / 'human_fast' stay headless no-ops with the same mode check so scripts run unchanged (render callbacks are
        not invoked: there is no EnvRenderer to hand them); 'rgb_array' returns the scene rasterised on the GPU as a
        (WINDOW_H, WINDOW_W, 3) uint8 numpy array, camera on the ego car."""
        assert mode in ['human', 'human_fast', 'rgb_array']
        if mode == 'rgb_array':
            return self.sim._batch.render(mode='rgb_array', envs=1)[0].cpu().numpy()
