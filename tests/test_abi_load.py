"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol the header
declares, and refuses to run without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from conftest import REPO


def test_library_exports_every_declared_symbol():
    from red_gym_b200 import _native as nat
    L = nat.load()
    header = open(os.path.join(REPO, "include", "f110_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)      # declarations only, not the prose
    declared = set(re.findall(r"\b(f110_[a-z_0-9]+)\s*\(", header))
    declared -= {"f110_sim"}
    assert declared == set(nat.EXPORTS), declared ^ set(nat.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name
    assert L.f110_abi_version() == nat.F110_ABI_VERSION


def test_config_struct_matches_header_layout():
    from red_gym_b200 import _native as nat
    # 12 int32 + 6 double + 1 uint64
    assert ctypes.sizeof(nat.Config) == 14 * 4 + 6 * 8 + 8
    assert ctypes.sizeof(nat.Buffers) == 16 * ctypes.sizeof(ctypes.c_void_p)


def test_invalid_arguments_are_rejected_before_touching_cuda():
    from red_gym_b200 import _native as nat
    L = nat.load()
    h = ctypes.c_void_p()
    cfg = nat.Config(abi_version=nat.F110_ABI_VERSION, n_envs=1, n_agents=1, n_beams=1080, integrator=7, ego_idx=0,
                     theta_dis=2000, timestep=0.01, fov=6.28, eps=1e-4, max_range=30.0, ttc_thresh=0.005, noise_std=0.01)
    assert L.f110_create(ctypes.byref(cfg), 0, ctypes.byref(h)) == nat.ERR_INVALID
    assert b"Integrator" in L.f110_last_error()
    cfg.integrator = 1
    cfg.n_agents = 99
    assert L.f110_create(ctypes.byref(cfg), 0, ctypes.byref(h)) == nat.ERR_INVALID
    cfg.n_agents = 1
    cfg.abi_version = 99
    assert L.f110_create(ctypes.byref(cfg), 0, ctypes.byref(h)) == nat.ERR_INVALID


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from red_gym_b200 import BatchF110Env, MAPS_DIR
    with pytest.raises(RuntimeError, match="no CUDA device"):
        BatchF110Env(num_envs=1, num_agents=1, map=os.path.join(MAPS_DIR, "example_map"))


def test_no_cpu_fallback_bitmap():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from red_gym_b200 import LidarBitmap
    with pytest.raises(RuntimeError, match="no CUDA device"):
        LidarBitmap(num_beams=1080)


def test_product_does_not_import_the_oracle():
    """The product path must never route through oracle/ (or any CPU fallback)."""
    for root in ("red_gym_b200", "f110_gym", "weap_util"):
        for dirpath, _, files in os.walk(os.path.join(REPO, root)):
            for f in files:
                if f.endswith((".py", ".cu", ".h")):
                    src = open(os.path.join(dirpath, f)).read()
                    assert "f110_oracle" not in src and "oracle/" not in src and "bitmap_oracle" not in src, os.path.join(dirpath, f)
