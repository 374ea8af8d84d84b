"""Host-side setup of the product vs the oracle's independent restatement (both follow the reference)."""
import os

import numpy as np
import pytest

import f110_oracle as fo
from conftest import MAPS
from red_gym_b200 import host


@pytest.mark.parametrize("fov", [2 * np.pi, 4.7])
@pytest.mark.parametrize("beams", [270, 1080, 2160])
def test_beam_tables(beams, fov):
    a = host.beam_tables(beams, fov, host.DEFAULT_PARAMS)
    b = fo.beam_tables(beams, fov, fo.DEFAULT_PARAMS)
    for x, y in zip(a, b):
        assert (x == y).all()


def test_trig_tables_have_endpoint():
    s, c = host.trig_tables(2000)
    so, co = fo.trig_tables(2000)
    assert (s == so).all() and (c == co).all()
    assert s.shape == (2000,) and abs(s[-1]) < 1e-15 and c[-1] == 1.0    # linspace includes 2*pi


@pytest.mark.parametrize("name", ["example_map", "berlin", "skirk"])
def test_map_edt(name):
    m = host.MapData(os.path.join(MAPS, name + ".yaml"), ".png")
    o = fo.load_map(os.path.join(MAPS, name + ".yaml"), ".png")
    assert (m.edt == o["dt"]).all() and m.height == o["h"] and m.width == o["w"]
    assert m.resolution == o["res"] and m.origin[0] == o["ox"] and m.origin[1] == o["oy"]
    assert m.edt.min() == 0.0 and m.edt.flags["C_CONTIGUOUS"]


def test_resolve_map_like_reference():
    assert host.resolve_map("berlin").endswith("maps/berlin.yaml")
    assert host.resolve_map(None).endswith("maps/vegas.yaml")
    assert host.resolve_map("/x/y/custom") == "/x/y/custom.yaml"


def test_noise_rows_are_the_numpy_stream():
    rng = np.random.default_rng(seed=12345)
    first = rng.normal(0., 0.01, size=1080)
    second = rng.normal(0., 0.01, size=1080)
    rows = host.noise_rows(12345, 0, 2, 1080)
    assert (rows[0] == first).all() and (rows[1] == second).all()
    assert (host.noise_rows(12345, 1, 1, 1080)[0] == second).all()


def test_params_vector_order_and_errors():
    v = host.params_vector(host.DEFAULT_PARAMS)
    assert v[0] == 1.0489 and v[16] == 0.31 and v[17] == 0.58 and v.shape == (18,)
    with pytest.raises(KeyError):
        host.params_vector({"mu": 1.0})


def test_start_rotation():
    r = host.start_rotation(np.array([0.3]))
    th = 0.3
    assert (r[0] == np.array([np.cos(-th), -np.sin(-th), np.sin(-th), np.cos(-th)])).all()


def test_noise_rows_drawn_ahead_are_the_streams_rows():
    """NoiseRowsAhead (parity-mode noise drawn by a helper thread): whatever the sizes asked for, the rows are the successive
    `default_rng(seed).normal(0, 0.01, B)` draws of the reference (base_classes.py:202, laser_models.py:451)."""
    from red_gym_b200.host import NoiseRowsAhead
    for seed, B in ((12345, 1080), (7, 33)):
        ahead = NoiseRowsAhead(seed, B)
        sizes = [128, 1, 300, 127, 2, 700]
        got = np.concatenate([ahead.take(n) for n in sizes], axis=0)
        ahead.close()
        rng = np.random.default_rng(seed=seed)
        want = np.stack([rng.normal(0., 0.01, size=B) for _ in range(sum(sizes))])
        assert got.shape == want.shape and (got == want).all()
