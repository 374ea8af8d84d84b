"""CPU: the bitmap oracle (oracle/bitmap_oracle.py, an OpenCV-free restatement of weap_util.lidar.lidar_to_bitmap)
against the golden images the unmodified reference function produced (tests/golden/bitmap.npz,
oracle/gen_bitmap_golden.py) and, where cv2 is importable, against cv2 itself on randomised primitives."""
import json
import os

import numpy as np
import pytest

import bitmap_oracle as bo
from conftest import GOLDEN


def _golden():
    d = np.load(os.path.join(GOLDEN, "bitmap.npz"))
    meta = json.loads(str(d["meta"]))
    scans = [d["scan_%d" % i] for i in range(int(d["n_scans"]))]
    out = []
    for k, m in enumerate(meta):
        shape = tuple(int(v) for v in d["shape_%d" % k])
        drawn = np.unpackbits(d["img_%d" % k])[:shape[0] * shape[1]].reshape(shape).astype(bool)
        kw = dict(m["kwargs"])
        if "output_image_dims" in kw:
            kw["output_image_dims"] = tuple(kw["output_image_dims"])
        bg, fg = (0, 255) if kw.get("bg_color", "white") == "black" else (255, 0)
        out.append((scans[m["scan"]], kw, np.where(drawn, fg, bg).astype(np.uint8)))
    return out


GOLD = _golden()


@pytest.mark.parametrize("k", range(len(GOLD)))
def test_oracle_matches_reference_images(k):
    scan, kw, want = GOLD[k]
    got = bo.lidar_to_bitmap(scan, **kw)
    assert got.dtype == np.uint8 and got.shape == want.shape
    assert (got == want).all(), int((got != want).sum())
    if kw["draw_mode"] == "FILL":      # the sort-free formulation the CUDA kernel uses gives the same pixels
        assert (bo.lidar_to_bitmap(scan, parity_fill=True, **kw) == want).all()


def test_channels_and_argument_checks():
    scan, kw, want = GOLD[0]
    rgb = bo.lidar_to_bitmap(scan, channels=3, **kw)
    assert rgb.shape == want.shape + (3,) and (rgb[..., 2] == want).all()
    rgba = bo.lidar_to_bitmap(scan, channels=4, **kw)
    assert rgba.shape == want.shape + (4,) and (rgba[..., 3] == 255).all()
    with pytest.raises(AssertionError):
        bo.lidar_to_bitmap(scan, draw_mode="SOLID")
    with pytest.raises(AssertionError):
        bo.lidar_to_bitmap(scan, target_beam_count=len(scan))
    with pytest.raises(ValueError):
        bo.lidar_to_bitmap(scan, scaling_factor=None)


def test_primitives_against_cv2_when_available():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    W = H = 64
    for _ in range(1500):
        p1, p2 = rng.integers(-40, 104, 2), rng.integers(-40, 104, 2)
        img = np.zeros((H, W), np.uint8)
        cv2.line(img, tuple(int(v) for v in p1), tuple(int(v) for v in p2), 255, 1)
        m = np.zeros((H, W), bool)
        for (x, y) in bo.line_pixels(W, H, p1, p2):
            m[y, x] = True
        assert (m == (img > 0)).all(), (p1, p2)
    for _ in range(1500):
        pts = rng.integers(-30, 94, (int(rng.integers(3, 12)), 2)).astype(np.int32)
        img = np.zeros((H, W), np.uint8)
        cv2.fillPoly(img, [pts], 255)
        assert (bo.fill_poly_mask(W, H, pts) == (img > 0)).all(), pts.tolist()
        m = bo.polylines_mask(W, H, pts)
        bo.fill_rows_parity(W, H, bo.poly_edges(W, H, pts), m)
        assert (m == (img > 0)).all(), pts.tolist()
