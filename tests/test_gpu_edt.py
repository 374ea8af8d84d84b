"""GPU parity of the device EDT (csrc/f110_edt.cu) against scipy.ndimage.distance_transform_edt, the third-party
routine the reference's get_dt calls (laser_models.py:40-53): bit-identical on every bundled map and on random images,
and the same scans through f110_set_map_image as through the host EDT."""
import os

import numpy as np
import pytest

from conftest import MAPS

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["example_map", "berlin", "skirk", "vegas", "stata_basement"])
def test_bundled_maps_bit_identical_to_scipy(name):
    from red_gym_b200 import host
    m = host.MapData(os.path.join(MAPS, name + ".yaml"), ".png")
    got = host.gpu_edt(m.free, m.resolution).cpu().numpy()
    assert got.shape == m.edt.shape and (got == m.edt).all(), float(np.abs(got - m.edt).max())


def test_random_images_and_edge_cases():
    from scipy.ndimage import distance_transform_edt
    from red_gym_b200 import host
    rng = np.random.default_rng(4)
    for (h, w, p) in ((1, 1, 0.5), (1, 300, 0.02), (300, 1, 0.02), (37, 911, 0.001), (700, 700, 0.0002), (513, 255, 0.3), (2300, 2100, 1e-6)):
        free = rng.uniform(size=(h, w)) > p
        free.flat[rng.integers(0, free.size)] = False          # at least one obstacle (scipy's answer is undefined without)
        want = 0.05 * distance_transform_edt(free)
        got = host.gpu_edt(free, 0.05).cpu().numpy()
        assert (got == want).all(), (h, w, p, float(np.abs(got - want).max()))


def test_env_with_device_edt_matches_host_edt():
    import torch
    from red_gym_b200 import BatchF110Env, host
    from red_gym_b200.host import raceline_poses
    poses = raceline_poses(500)
    kw = dict(num_envs=1, num_agents=1, map=os.path.join(MAPS, "example_map"), noise="off")
    for dtype in ("fp64", "fp32"):
        a = BatchF110Env(dtype=dtype, edt="scipy", **kw)     # get_dt on the host, like the reference
        b = BatchF110Env(dtype=dtype, **kw)                  # default: device EDT
        assert b.map.edt is None and a.map.edt is not None
        assert (a.raycast(poses) == b.raycast(poses)).all()
        b.update_map(os.path.join(MAPS, "berlin.yaml"), ".png")            # default: device EDT
        a.update_map(os.path.join(MAPS, "berlin.yaml"), ".png", edt="scipy")
        p2 = torch.tensor([[0.0, 0.0, 0.3], [1.0, 0.5, 2.0]], dtype=torch.float64)
        assert (a.raycast(p2) == b.raycast(p2)).all()
