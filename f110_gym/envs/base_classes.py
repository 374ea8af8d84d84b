"""Drop-in for the reference's base_classes (base_classes.py:40-623): `Integrator`, a `Simulator` whose step runs on the
GPU, and read-only `RaceCar` views of its agents."""
from enum import Enum

import numpy as np

from red_gym_b200 import host
from red_gym_b200.batch_env import BatchF110Env


class Integrator(Enum):
    RK4 = 1
    Euler = 2


class RaceCar(object):
    """What `sim.agents[i]` exposes in the reference (base_classes.py:44-156), as a read-only view of agent i of the
    device-side simulator: `state` (7,) = [x, y, steer_angle, v, yaw, yaw_rate, slip] (:95-96), `in_collision` (:112),
    `steer_buffer` (:106), the construction data (`params`, `seed`, `is_ego`, `time_step`, `num_beams`, `fov`, `integrator`,
    `ttc_thresh` :113) and the per-beam tables that are class statics there (`scan_angles`, `cosines`, `side_distances`
    :63-66,116-156).  Values are fetched from the GPU on access.  The per-agent methods (update_pose, check_ttc, ...) are
    not offered: the whole tick of all agents is one library call (`Simulator.step`); `accel` / `steer_angle_vel` (the
    last PID output, :99-100) are internal to the dynamics kernel and not kept."""

    def __init__(self, sim, index):
        self._sim, self._i = sim, index
        self.params = sim.params
        self.seed = sim.seed
        self.is_ego = index == sim.ego_idx
        self.time_step = sim.time_step
        self.num_beams = 1080
        self.fov = sim.fov
        self.integrator = sim.integrator
        self.ttc_thresh = 0.005
        self.steer_buffer_size = 2

    def _tables(self):
        s = self._sim
        if s._tables is None:
            s._tables = host.beam_tables(self.num_beams, s.fov, s.params)
        return s._tables

    @property
    def scan_angles(self):
        return self._tables()[0]

    @property
    def cosines(self):
        return self._tables()[1]

    @property
    def side_distances(self):
        return self._tables()[2]

    @property
    def state(self):
        b = self._sim._need_batch()
        return b.state[:, 0, self._i].cpu().numpy()

    @property
    def in_collision(self):
        return bool(self._sim._need_batch().in_collision[0, self._i].item())

    @property
    def scan_simulator(self):
        return self._sim._scan_simulator()


class Simulator(object):
    """Same constructor and methods as the reference Simulator (base_classes.py:445-623) for one env.
    The whole tick (dynamics, scans, iTTC, GJK, opponent ray-cast and the env's lap bookkeeping) is one
    C-ABI call; results come back through pinned host buffers.

    `ego_idx` is what the observation reports (the reference's F110Env never forwards its own ego_idx to the Simulator,
    f110_env.py:190, so obs['ego_idx'] is 0 there whatever the env was given); `env_ego_idx` is the env-level ego used
    for `done` and the lap frame (f110_env.py:292-299, 318-329), which live in the same library call here."""

    def __init__(self, params, num_agents, seed, fov, time_step=0.01, ego_idx=0, integrator=Integrator.RK4,
                 map_path=None, map_ext='.png', device=0, env_ego_idx=None):
        self.num_agents = num_agents
        self.seed = seed
        self.time_step = time_step
        self.ego_idx = ego_idx
        self._env_ego_idx = ego_idx if env_ego_idx is None else env_ego_idx
        self.params = params
        self.fov = fov
        self.integrator = integrator
        self.agent_poses = np.empty((self.num_agents, 3))
        self.collisions = np.zeros((self.num_agents, ))
        self._collision_idx = -1 * np.ones((self.num_agents, ))
        self._device = device
        self._batch = None
        self._tables = None
        self._scan_sim = None
        self._map = None              # (yaml path, image extension) once set_map was called
        self.agents = [RaceCar(self, i) for i in range(num_agents)]
        if map_path is not None:
            self.set_map(map_path, map_ext)

    def _need_batch(self):
        if self._batch is None:
            raise ValueError('Map is not set for scan simulator.')
        return self._batch

    def _scan_simulator(self):
        from f110_gym.envs.laser_models import ScanSimulator2D
        if self._scan_sim is None and self._map is not None:
            self._scan_sim = ScanSimulator2D(1080, self.fov, device=self._device)
            self._scan_sim.set_map(*self._map)
        return self._scan_sim

    def _make(self, map_yaml, map_ext):
        import torch
        kw = dict(seed=self.seed, params=self.params, num_agents=self.num_agents, timestep=self.time_step,
                  ego_idx=self._env_ego_idx, integrator=self.integrator, fov=self.fov, map_ext=map_ext)
        kw['map'] = map_yaml[:-5] if map_yaml.endswith('.yaml') else map_yaml
        self._batch = BatchF110Env(num_envs=1, device=self._device, dtype='fp64', noise='numpy', **kw)
        A = self.num_agents
        self._h_act = torch.zeros((1, A, 2), dtype=torch.float64).pin_memory()
        self._h_obs = torch.zeros((10, A), dtype=torch.float64).pin_memory()
        self._h_scans = torch.zeros((A, 1080), dtype=torch.float64).pin_memory()
        self._h_done = torch.zeros((1,), dtype=torch.uint8).pin_memory()
        # numpy views and raw addresses of the pinned buffers, made once: a single-env step is ~35 us of GPU work
        self._np_act, self._np_obs, self._np_scans = self._h_act.numpy(), self._h_obs.numpy(), self._h_scans.numpy()
        self._np_done = self._h_done.numpy()
        self._ptrs = (self._h_act.data_ptr(), self._h_obs.data_ptr(), self._h_scans.data_ptr(), self._h_done.data_ptr())

    def set_map(self, map_path, map_ext):
        self._map = (map_path, map_ext)
        self._scan_sim = None
        if self._batch is None:
            self._make(map_path, map_ext)
        else:
            self._batch.update_map(map_path, map_ext)

    def update_params(self, params, agent_idx=-1):
        if agent_idx >= self.num_agents:
            raise IndexError('Index given is out of bounds for list of agents.')
        self._need_batch().update_params(params, index=agent_idx)

    def _collect(self):
        o = self._np_obs.copy()                       # (10, A): one copy, then plain views / scalars of it
        A = self.num_agents
        self.agent_poses = o[[0, 1, 4]].T
        self.collisions = o[7]
        self._collision_idx = None if A > 1 else -1 * np.ones((A,))      # A > 1: fetched from the device on access
        scans = self._np_scans
        observations = {'ego_idx': self.ego_idx,
                        'scans': [scans[i].copy() for i in range(A)],
                        'poses_x': [np.float64(o[0, i]) for i in range(A)],
                        'poses_y': [np.float64(o[1, i]) for i in range(A)],
                        'poses_theta': [np.float64(o[4, i]) for i in range(A)],
                        'linear_vels_x': [np.float64(o[3, i]) for i in range(A)],
                        'linear_vels_y': [0. for _ in range(A)],
                        'ang_vels_z': [np.float64(o[5, i]) for i in range(A)],
                        'collisions': self.collisions}
        self.lap_times = o[8]
        self.lap_counts = o[9]
        self.done = bool(self._np_done[0])
        return observations

    @property
    def collision_idx(self):
        """collision_multiple's partner index per agent (collision_models.py:197), -1 if none."""
        if self._collision_idx is None:
            self._collision_idx = self._batch.collision_idx.cpu().numpy()[0]
        return self._collision_idx

    def step(self, control_inputs):
        """Simulator.step (base_classes.py:546-605): one tick of all agents -> observation dict."""
        b = self._need_batch()
        self._np_act[0] = np.asarray(control_inputs, dtype=np.float64).reshape(self.num_agents, 2)
        b.step_host_ptrs(*self._ptrs)
        return self._collect()

    def reset(self, poses):
        """Simulator.reset (base_classes.py:607-623): resets the agents to `poses` and returns None -- no step is taken
        (the zero-action step after a reset belongs to F110Env.reset, f110_env.py:335-336)."""
        poses = np.asarray(poses, dtype=np.float64)
        if poses.shape[0] != self.num_agents:
            raise ValueError('Number of poses for reset does not match number of agents.')
        self._need_batch().reset(poses[None], zero_action_step=False)
        self.collisions = np.zeros((self.num_agents, ))
        self._collision_idx = -1 * np.ones((self.num_agents, ))
