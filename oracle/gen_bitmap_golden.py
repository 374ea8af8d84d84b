"""Golden images of the scan -> bitmap rasteriser: runs the UNMODIFIED reference function
weap_util.lidar.lidar_to_bitmap (/root/reference/weap_util/weap_util/lidar.py:114, OpenCV drawing underneath) in
THIS container and writes tests/golden/bitmap.npz.  Test infrastructure only.

    python oracle/gen_bitmap_golden.py

Scans: real lidar scans of the reference simulator (tests/golden/scans_example_map_2pi_b1080.npz, made by
oracle/gen_golden.py) plus synthetic ones that push polygon vertices far outside the image (ranges up to 30 m at
25 px/m), zero ranges, and a float32 copy (numpy then multiplies scaling_factor * data in float32).
Images are stored bit-packed (np.packbits) with their keyword arguments; cv2 version recorded.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
GOLDEN = os.path.join(REPO, "tests", "golden")


def cases():
    real = np.load(os.path.join(GOLDEN, "scans_example_map_2pi_b1080.npz"))["scans"]
    rng = np.random.default_rng(2025)
    scans = [real[i] for i in (0, 3, 7, 11)]
    B = 1080
    base = rng.uniform(0.5, 30, 8)
    wavy = np.clip(np.interp(np.linspace(0, 8, B), np.arange(9), np.append(base, base[0])) + rng.normal(0, 0.3, B), 0, 30)
    spiky = wavy.copy(); spiky[rng.integers(0, B, 80)] = 30.0
    zeros = np.zeros(B); zeros[100:300] = 12.0
    scans += [wavy, spiky, zeros, real[5].astype(np.float32)]
    kws = [dict(draw_mode="FILL", bg_color="black"),                                        # src/SAL.py:76-77
           dict(draw_mode="POLYGON"),                                                       # the function's defaults
           dict(draw_mode="RAYS", bg_color="black", target_beam_count=50, fov=4.7),         # examples/lidar_example.py:104
           dict(draw_mode="FILL", scaling_factor=25, winding_dir="CW", draw_center=False),
           dict(draw_mode="FILL", max_scan_radius=30.0, output_image_dims=(128, 128), target_beam_count=300),
           dict(draw_mode="POLYGON", scaling_factor=25, starting_angle=0.3, fov=4.7),
           dict(draw_mode="RAYS", scaling_factor=4, target_beam_count=200)]
    out = []
    for si, s in enumerate(scans):
        for ki, kw in enumerate(kws):
            if (si + ki) % 2 == 0 or ki == 0:
                out.append((si, kw))
    return scans, out


def main():
    sys.path.insert(0, "/root/reference/weap_util")
    import cv2
    from weap_util.lidar import lidar_to_bitmap
    scans, todo = cases()
    data = {"n_scans": len(scans), "cv2_version": cv2.__version__}
    for i, s in enumerate(scans):
        data["scan_%d" % i] = s
    meta = []
    for k, (si, kw) in enumerate(todo):
        img = lidar_to_bitmap(scans[si], **kw)
        fg = 255 if kw.get("bg_color", "white") == "black" else 0
        assert set(np.unique(img)) <= {0, 255}
        data["img_%d" % k] = np.packbits(img == fg)
        data["shape_%d" % k] = np.array(img.shape)
        meta.append({"scan": si, "kwargs": kw})
    # channels (lidar.py:147-154)
    data["rgb"] = np.packbits(lidar_to_bitmap(scans[0], draw_mode="FILL", bg_color="black", channels=3)[..., 1] == 255)
    data["rgba_alpha_min"] = lidar_to_bitmap(scans[0], channels=4)[..., 3].min()
    data["meta"] = json.dumps(meta)
    np.savez_compressed(os.path.join(GOLDEN, "bitmap.npz"), **data)
    print("wrote", len(todo), "images, cv2", cv2.__version__)


if __name__ == "__main__":
    main()
