"""GPU-backed stand-in for the reference's laser_models.ScanSimulator2D (laser_models.py:348-457)."""
import numpy as np

from red_gym_b200 import host
from red_gym_b200.batch_env import BatchF110Env


class ScanSimulator2D(object):
    """Same constructor, `set_map`, `scan(pose, rng, std_dev)` and `get_increment` as the reference
    (laser_models.py:360-457); the ray marching runs on the GPU (f110_raycast).  eps, theta_dis and max_range are
    passed on to the library like any other value of f110_config."""

    def __init__(self, num_beams, fov, eps=0.0001, theta_dis=2000, max_range=30.0, device=0):
        self.num_beams = num_beams
        self.fov = fov
        self.eps = eps
        self.theta_dis = theta_dis
        self.max_range = max_range
        self.angle_increment = self.fov / (self.num_beams - 1)
        self.theta_index_increment = theta_dis * self.angle_increment / (2. * np.pi)
        self.map_height = None
        self.map_width = None
        self.map_resolution = None
        self.dt = None
        self._device = device
        self._env = None
        self.sines, self.cosines = host.trig_tables(theta_dis)

    def set_map(self, map_path, map_ext):
        m = host.MapData(map_path, map_ext)
        self.map_height, self.map_width, self.map_resolution = m.height, m.width, m.resolution
        self.origin = list(m.origin)
        self.orig_x, self.orig_y = m.origin[0], m.origin[1]
        self.orig_c, self.orig_s = m.orig_c, m.orig_s
        self.dt = m.edt
        self._env = BatchF110Env(num_envs=1, num_agents=1, num_beams=self.num_beams, fov=self.fov, noise='off',
                                 device=self._device, map_data=m, eps=self.eps, theta_dis=self.theta_dis, max_range=self.max_range)
        return True

    def scan(self, pose, rng, std_dev=0.01):
        if self.map_height is None:
            raise ValueError('Map is not set for scan simulator.')
        scan = self._env.raycast(np.asarray(pose, dtype=np.float64).reshape(1, 3)).cpu().numpy()[0]
        if rng is not None:
            scan += rng.normal(0., std_dev, size=self.num_beams)
        return scan

    def scan_batch(self, poses):
        """(n, 3) poses -> (n, num_beams) CUDA tensor, no noise."""
        if self.map_height is None:
            raise ValueError('Map is not set for scan simulator.')
        return self._env.raycast(poses)

    def get_increment(self):
        return self.angle_increment
