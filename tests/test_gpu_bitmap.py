"""GPU parity of the scan -> bitmap rasteriser (csrc/f110_bitmap.cu through the C ABI): byte-exact against the golden
images of the unmodified reference function, against the oracle on randomised batches in every mode, and on the
simulator's own scans (fp64 and fp32)."""
import os

import numpy as np
import pytest

import bitmap_oracle as bo
from conftest import MAPS
from test_bitmap_oracle import GOLD

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k", range(len(GOLD)))
def test_kernel_matches_reference_images(k):
    import torch
    from red_gym_b200 import LidarBitmap
    scan, kw, want = GOLD[k]
    dt = "fp32" if scan.dtype == np.float32 else "fp64"
    bm = LidarBitmap(num_beams=len(scan), scan_dtype=dt, **kw)
    got = bm(torch.from_numpy(scan)[None].cuda())[0].cpu().numpy()
    assert got.shape == want.shape and (got == want).all(), int((got != want).sum())
    packed = bm(torch.from_numpy(scan)[None].cuda(), packed=True)[0].cpu().numpy().view(np.uint32)
    bits = ((packed[:, :, None] >> np.arange(32, dtype=np.uint32)) & 1).reshape(want.shape[0], -1)[:, :want.shape[1]].astype(bool)
    fg = 255 if kw.get("bg_color", "white") == "black" else 0
    assert (bits == (want == fg)).all()


@pytest.mark.parametrize("mode", ["FILL", "POLYGON", "RAYS"])
@pytest.mark.parametrize("dtype", ["fp64", "fp32"])
def test_random_batches_vs_oracle(mode, dtype):
    import torch
    from red_gym_b200 import LidarBitmap
    rng = np.random.default_rng(17)
    n, B = 48, 1080
    base = rng.uniform(0.3, 30, (n, 12))
    scans = np.stack([np.interp(np.linspace(0, 12, B), np.arange(13), np.append(b, b[0])) for b in base])
    scans = np.clip(scans + rng.normal(0, 0.25, scans.shape), 0, 30)
    scans[::5, ::37] = 30.0
    scans[3] = 0.0
    if dtype == "fp32":
        scans = scans.astype(np.float32)
    for kw in (dict(scaling_factor=10, bg_color="black"), dict(scaling_factor=25, winding_dir="CW", fov=4.7),
               dict(max_scan_radius=20.0, output_image_dims=(96, 160), target_beam_count=150)):
        bm = LidarBitmap(num_beams=B, draw_mode=mode, scan_dtype=dtype, **kw)
        got = bm(torch.from_numpy(scans).cuda()).cpu().numpy()
        for i in range(0, n, 3 if mode != "RAYS" else 8):
            want = bo.lidar_to_bitmap(scans[i], draw_mode=mode, **kw)
            assert (got[i] == want).all(), (mode, dtype, kw, i, int((got[i] != want).sum()))


def test_simulator_scans_and_dropin_function():
    """The rasteriser on the env's own scans (the call src/SAL.py makes every step), through the weap_util import path."""
    import torch
    from red_gym_b200 import BatchF110Env, LidarBitmap
    from red_gym_b200.host import raceline_poses
    from weap_util.lidar import lidar_to_bitmap
    N = 64
    for dtype, noise in (("fp64", "numpy"), ("fp32", "philox")):
        env = BatchF110Env(num_envs=N, num_agents=1, map=os.path.join(MAPS, "example_map"), dtype=dtype, noise=noise)
        obs, _, _, _ = env.reset(raceline_poses(N)[:, None, :])
        bm = LidarBitmap(num_beams=1080, draw_mode="FILL", bg_color="black", scan_dtype=dtype)
        imgs = bm(obs["scans"].view(N, 1080)).cpu().numpy()
        scans = obs["scans"].view(N, 1080).cpu().numpy()
        for i in range(0, N, 7):
            assert (imgs[i] == bo.lidar_to_bitmap(scans[i], draw_mode="FILL", bg_color="black")).all()
        one = lidar_to_bitmap(scans[5], output_image_dims=(256, 256), bg_color="black", draw_mode="FILL")
        assert one.dtype == np.uint8 and (one == imgs[5]).all()
        rgba = lidar_to_bitmap(scans[5], bg_color="black", draw_mode="FILL", channels=4)
        assert rgba.shape == (256, 256, 4) and (rgba[..., 0] == imgs[5]).all() and (rgba[..., 3] == 255).all()
    with pytest.raises(AssertionError):
        LidarBitmap(num_beams=1080, draw_mode="SOLID")
    with pytest.raises(AssertionError):
        LidarBitmap(num_beams=1080, target_beam_count=1080)
