"""CPU oracle of the headless renderer.  TEST INFRASTRUCTURE ONLY.

The reference draws with pyglet/OpenGL (gym/f110_gym/envs/rendering.py); GL is unavailable on this box, so there is no
reference output to pin against: PARITY UNPINNED.  This file restates the scene and the projection of the reference in
numpy and is the definition the CUDA rasteriser is checked against:
  * clear colour (9, 32, 87)                                         rendering.py:72
  * one point per map pixel whose value is exactly 0, at 50 * (col * res + ox, row * res + oy), colour (183, 193, 222)
                                                                     rendering.py:134-153 (image flipped top-bottom)
  * one quad per car, 50 * get_vertices(pose, 0.58, 0.31); ego (172, 97, 185), others (99, 52, 94), drawn over the map
                                                                     rendering.py:46-47, 316-334; collision_models.py:219-260
  * orthographic camera glOrtho(left, right, bottom, top)            rendering.py:290; default window 1000 x 800 with
    zoom 1.2 (rendering.py:74-80,178-184), or following the ego car like examples/waypoint_follow.py:259-269
    (car bounding box +- 800)
Rasterisation rules (GL's, for aliased primitives): a point lights the pixel that contains it; a quad lights the pixels
whose centre lies inside it (edges inclusive here).  Text (lap label, fps) is left out.  Row 0 of the returned image is the
TOP of the window.
"""
import numpy as np

CAR_LENGTH, CAR_WIDTH = 0.58, 0.31
CLEAR = (9, 32, 87)
MAP_COLOR = (183, 193, 222)
EGO_COLOR = (172, 97, 185)
OPP_COLOR = (99, 52, 94)


def get_vertices(pose, length=CAR_LENGTH, width=CAR_WIDTH):
    """collision_models.get_vertices: corners rl, rr, fr, fl."""
    x, y, th = float(pose[0]), float(pose[1]), float(pose[2])
    c, s = np.cos(th), np.sin(th)
    out = []
    for (lx, ly) in ((-length / 2, width / 2), (-length / 2, -width / 2), (length / 2, -width / 2), (length / 2, width / 2)):
        out.append((c * lx + -s * ly + x, s * lx + c * ly + y))
    return np.array(out)


def default_camera(width=1000, height=800, zoom=1.2):
    return (-zoom * width / 2, zoom * width / 2, -zoom * height / 2, zoom * height / 2)


def follow_camera(ego_pose, margin=800.0):
    v = 50.0 * get_vertices(ego_pose)
    return (v[:, 0].min() - margin, v[:, 0].max() + margin, v[:, 1].min() - margin, v[:, 1].max() + margin)


def render(obstacle, res, ox, oy, poses, ego_idx=0, camera=None, width=1000, height=800):
    """obstacle: (h, w) bool, True where the (flipped) map image is exactly 0; poses: (A, 3).  -> (height, width, 3) uint8."""
    left, right, bottom, top = camera if camera is not None else default_camera(width, height)
    img = np.empty((height, width, 3), dtype=np.uint8)
    img[:] = CLEAR
    rows, cols = np.nonzero(obstacle)
    xw = (50.0 * (cols * res + ox) - left) / (right - left) * width
    yw = (50.0 * (rows * res + oy) - bottom) / (top - bottom) * height
    px, py = np.floor(xw).astype(np.int64), np.floor(yw).astype(np.int64)
    ok = (px >= 0) & (px < width) & (py >= 0) & (py < height)
    img[height - 1 - py[ok], px[ok]] = MAP_COLOR
    X = left + (np.arange(width) + 0.5) * ((right - left) / width)
    Y = bottom + (np.arange(height) + 0.5) * ((top - bottom) / height)
    for a in range(len(poses)):
        v = 50.0 * get_vertices(poses[a])
        inside_pos = np.ones((height, width), dtype=bool)
        inside_neg = np.ones((height, width), dtype=bool)
        for k in range(4):
            x0, y0 = v[k]
            x1, y1 = v[(k + 1) % 4]
            cr = (x1 - x0) * (Y[:, None] - y0) - (y1 - y0) * (X[None, :] - x0)
            inside_pos &= cr >= 0
            inside_neg &= cr <= 0
        m = (inside_pos | inside_neg)[::-1]
        img[m] = EGO_COLOR if a == ego_idx else OPP_COLOR
    return img
