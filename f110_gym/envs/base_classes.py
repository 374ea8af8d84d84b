"""Drop-in for the reference's base_classes (base_classes.py:40-623): `Integrator` and a
`Simulator` whose step runs on the GPU."""
from enum import Enum

import numpy as np

from red_gym_b200.batch_env import BatchF110Env


class Integrator(Enum):
    RK4 = 1
    Euler = 2


class Simulator(object):
    """Same constructor and methods as the reference Simulator (base_classes.py:445-623) for one env.
    The whole tick (dynamics, scans, iTTC, GJK, opponent ray-cast and the env's lap bookkeeping) is one
    C-ABI call; results come back through pinned host buffers."""

    def __init__(self, params, num_agents, seed, fov, time_step=0.01, ego_idx=0, integrator=Integrator.RK4,
                 map_path=None, map_ext='.png', device=0):
        self.num_agents = num_agents
        self.seed = seed
        self.time_step = time_step
        self.ego_idx = ego_idx
        self.params = params
        self.fov = fov
        self.integrator = integrator
        self.agent_poses = np.empty((self.num_agents, 3))
        self.collisions = np.zeros((self.num_agents, ))
        self._collision_idx = -1 * np.ones((self.num_agents, ))
        self._device = device
        self._batch = None
        if map_path is not None:
            self.set_map(map_path, map_ext)

    def _make(self, map_yaml, map_ext):
        import torch
        kw = dict(seed=self.seed, params=self.params, num_agents=self.num_agents, timestep=self.time_step,
                  ego_idx=self.ego_idx, integrator=self.integrator, fov=self.fov, map_ext=map_ext)
        kw['map'] = map_yaml[:-5] if map_yaml.endswith('.yaml') else map_yaml
        self._batch = BatchF110Env(num_envs=1, device=self._device, dtype='fp64', noise='numpy', **kw)
        A = self.num_agents
        self._h_act = torch.zeros((1, A, 2), dtype=torch.float64).pin_memory()
        self._h_obs = torch.zeros((10, A), dtype=torch.float64).pin_memory()
        self._h_scans = torch.zeros((A, 1080), dtype=torch.float64).pin_memory()
        self._h_done = torch.zeros((1,), dtype=torch.uint8).pin_memory()

    def set_map(self, map_path, map_ext):
        if self._batch is None:
            self._make(map_path, map_ext)
        else:
            self._batch.update_map(map_path, map_ext)

    def update_params(self, params, agent_idx=-1):
        if agent_idx >= self.num_agents:
            raise IndexError('Index given is out of bounds for list of agents.')
        self._batch.update_params(params, index=agent_idx)

    def _collect(self):
        o = self._h_obs.numpy()
        A = self.num_agents
        self.agent_poses = np.stack([o[0], o[1], o[4]], axis=1)
        self.collisions = o[7].copy()
        self._collision_idx = None if A > 1 else -1 * np.ones((A,))      # A > 1: fetched from the device on access
        scans = self._h_scans.numpy()
        observations = {'ego_idx': self.ego_idx,
                        'scans': [scans[i].copy() for i in range(A)],
                        'poses_x': [np.float64(o[0, i]) for i in range(A)],
                        'poses_y': [np.float64(o[1, i]) for i in range(A)],
                        'poses_theta': [np.float64(o[4, i]) for i in range(A)],
                        'linear_vels_x': [np.float64(o[3, i]) for i in range(A)],
                        'linear_vels_y': [0. for _ in range(A)],
                        'ang_vels_z': [np.float64(o[5, i]) for i in range(A)],
                        'collisions': self.collisions}
        self.lap_times = o[8].copy()
        self.lap_counts = o[9].copy()
        self.done = bool(self._h_done[0])
        return observations

    @property
    def collision_idx(self):
        """collision_multiple's partner index per agent (collision_models.py:197), -1 if none."""
        if self._collision_idx is None:
            self._collision_idx = self._batch.collision_idx.cpu().numpy()[0]
        return self._collision_idx

    def step(self, control_inputs):
        if self._batch is None:
            raise ValueError('Map is not set for scan simulator.')
        self._h_act.numpy()[0] = np.asarray(control_inputs, dtype=np.float64).reshape(self.num_agents, 2)
        self._batch.step_host(self._h_act, self._h_obs, self._h_scans, self._h_done)
        return self._collect()

    def reset(self, poses):
        poses = np.asarray(poses, dtype=np.float64)
        if poses.shape[0] != self.num_agents:
            raise ValueError('Number of poses for reset does not match number of agents.')
        if self._batch is None:
            raise ValueError('Map is not set for scan simulator.')
        self._batch.reset(poses[None])
        self._h_obs.copy_(self._batch.obs_block.reshape(10, self.num_agents))
        self._h_scans.copy_(self._batch.scans[0])
        self._h_done.copy_(self._batch.done)
        return self._collect()
