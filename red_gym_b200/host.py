"""Host-side setup for the batched simulator: map -> EDT, lookup tables, noise rows, params.

Setup-time work the reference also does on the host (ScanSimulator2D.__init__/set_map,
RaceCar.__init__); the arrays are uploaded once through the C ABI.  Values must be
bit-identical to the reference's so that cell indices and iTTC flags match, hence numpy
float64 throughout and the same formulas (cited per function).
"""
import os

import numpy as np

from ._native import PARAM_KEYS

MAPS_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "maps")

# reference default vehicle (f110_env.py:128)
DEFAULT_PARAMS = {"mu": 1.0489, "C_Sf": 4.718, "C_Sr": 5.4562, "lf": 0.15875, "lr": 0.17145, "h": 0.074,
                  "m": 3.74, "I": 0.04712, "s_min": -0.4189, "s_max": 0.4189, "sv_min": -3.2, "sv_max": 3.2,
                  "v_switch": 7.319, "a_max": 9.51, "v_min": -5.0, "v_max": 20.0, "width": 0.31, "length": 0.58}


def resolve_map(map_name):
    """kwargs['map'] -> yaml path, as F110Env.__init__ does (f110_env.py:106-118): bundled names map to
    the packaged yaml, anything else is '<map>.yaml'."""
    if map_name is None:
        return os.path.join(MAPS_DIR, "vegas.yaml")
    if map_name in ("berlin", "skirk", "levine"):
        return os.path.join(MAPS_DIR, map_name + ".yaml")
    return map_name + ".yaml"


class MapData(object):
    """Occupancy image -> metric EDT (ScanSimulator2D.set_map, laser_models.py:383-427)."""

    def __init__(self, map_path, map_ext, edt="scipy", origin_yaw=None):
        """edt='scipy': get_dt on the host like the reference; edt='gpu': keep only the thresholded image, the library
        builds the (bit-identical) EDT on the device when the map is uploaded (f110_set_map_image).
        origin_yaw: replaces the yaml's origin[2] (the map-frame rotation of xy_2_rc, laser_models.py:75-76,421-422)."""
        import yaml
        from PIL import Image
        img_path = os.path.splitext(map_path)[0] + map_ext
        occ = np.asarray(Image.open(img_path).transpose(Image.FLIP_TOP_BOTTOM), dtype=np.float64)
        free = occ > 128.0                                   # <=128 -> obstacle, >128 -> free (:403-404)
        self.free = np.ascontiguousarray(free, dtype=np.uint8)
        self.black = np.ascontiguousarray(occ == 0.0, dtype=np.uint8)     # the pixels EnvRenderer.update_map draws (rendering.py:150)
        with open(map_path, "r") as stream:
            meta = yaml.safe_load(stream)
        self.resolution = float(meta["resolution"])
        origin = list(meta["origin"])
        if origin_yaw is not None:
            origin[2] = float(origin_yaw)
        self.origin = (float(origin[0]), float(origin[1]), float(origin[2]))
        self.orig_c = float(np.cos(origin[2]))
        self.orig_s = float(np.sin(origin[2]))
        self.height, self.width = occ.shape[0], occ.shape[1]
        # get_dt (laser_models.py:40-53): exact Euclidean distance to the nearest obstacle pixel, metres
        if edt == "scipy":
            from scipy.ndimage import distance_transform_edt
            self.edt = np.ascontiguousarray(self.resolution * distance_transform_edt(free))
        elif edt == "gpu":
            self.edt = None
        else:
            raise ValueError("edt must be 'scipy' or 'gpu'")
        self.path = map_path
        self.ext = map_ext

    @classmethod
    def from_free_mask(cls, free, resolution, origin=(0.0, 0.0, 0.0)):
        """A map given as its thresholded image (non-zero = free, row 0 = bottom); the EDT is built on the device."""
        self = cls.__new__(cls)
        self.free = np.ascontiguousarray(np.asarray(free) != 0, dtype=np.uint8)
        self.edt = None
        self.height, self.width = self.free.shape
        self.resolution = float(resolution)
        self.origin = tuple(float(o) for o in origin)
        self.orig_c = float(np.cos(origin[2]))
        self.orig_s = float(np.sin(origin[2]))
        self.path = None
        self.ext = None
        return self

    @classmethod
    def from_arrays(cls, edt, resolution, origin=(0.0, 0.0, 0.0)):
        self = cls.__new__(cls)
        self.free = None
        self.edt = np.ascontiguousarray(edt, dtype=np.float64)
        self.height, self.width = self.edt.shape
        self.resolution = float(resolution)
        self.origin = tuple(float(o) for o in origin)
        self.orig_c = float(np.cos(origin[2]))
        self.orig_s = float(np.sin(origin[2]))
        self.path = None
        self.ext = None
        return self


def trig_tables(theta_dis=2000):
    """sin/cos of np.linspace(0, 2*pi, theta_dis) -- endpoint included (laser_models.py:379-381)."""
    theta = np.linspace(0.0, 2 * np.pi, num=theta_dis)
    return np.sin(theta), np.cos(theta)


def beam_tables(num_beams, fov, params):
    """Per-beam angle, cosine and lidar-to-hull distance (RaceCar.__init__, base_classes.py:120-156)."""
    inc = fov / (num_beams - 1)
    ang = -fov / 2. + np.arange(num_beams) * inc
    cosines = np.cos(ang)
    half_w = params["width"] / 2.
    half_l = (params["lf"] + params["lr"]) / 2.
    # each quadrant folds onto the first one: distance along the beam to the side and to the front/rear
    with np.errstate(divide="ignore", invalid="ignore"):
        q1 = (ang > 0) & (ang < np.pi / 2)
        q2 = (ang > 0) & ~(ang < np.pi / 2)
        q4 = ~(ang > 0) & (ang > -np.pi / 2)
        q3 = ~(ang > 0) & ~(ang > -np.pi / 2)
        to_side = np.empty(num_beams)
        to_fr = np.empty(num_beams)
        to_side[q1] = half_w / np.sin(ang[q1]);               to_fr[q1] = half_l / np.cos(ang[q1])
        to_side[q2] = half_w / np.cos(ang[q2] - np.pi / 2.);  to_fr[q2] = half_l / np.sin(ang[q2] - np.pi / 2.)
        to_side[q4] = half_w / np.sin(-ang[q4]);              to_fr[q4] = half_l / np.cos(-ang[q4])
        to_side[q3] = half_w / np.cos(-ang[q3] - np.pi / 2);  to_fr[q3] = half_l / np.sin(-ang[q3] - np.pi / 2)
    side = np.minimum(to_side, to_fr)
    return ang, cosines, side


def noise_rows(seed, first, count, num_beams, std_dev=0.01):
    """Rows [first, first+count) of the per-car noise stream: row k is the k-th
    `default_rng(seed).normal(0., std, size=B)` draw after a reset (base_classes.py:202, laser_models.py:451).
    PCG64 + ziggurat is sequential, so rows are regenerated from the start of the stream."""
    rng = np.random.default_rng(seed=seed)
    out = np.empty((first + count, num_beams))
    for k in range(first + count):
        out[k] = rng.normal(0., std_dev, size=num_beams)
    return out[first:]


def params_vector(params):
    missing = [k for k in PARAM_KEYS if k not in params]
    if missing:
        raise KeyError("vehicle params missing keys: %s" % missing)
    return np.array([float(params[k]) for k in PARAM_KEYS], dtype=np.float64)


def start_rotation(theta_ego):
    """Rotation by -theta_ego, elementwise like f110_env.py:329; theta_ego: (N,) -> (N, 4) row-major."""
    t = np.asarray(theta_ego, dtype=np.float64)
    return np.ascontiguousarray(np.stack([np.cos(-t), -np.sin(-t), np.sin(-t), np.cos(-t)], axis=-1))


def raceline_poses(n, csv_path=None, lateral=0.0, seed=0):
    """n start poses spread along the example_map raceline (x, y, psi + pi/2) -- SURVEY 8d config 2."""
    csv_path = csv_path or os.path.join(MAPS_DIR, "example_waypoints.csv")
    wp = np.loadtxt(csv_path, delimiter=";", skiprows=3)
    idx = (np.arange(n) * wp.shape[0]) // n
    poses = np.stack([wp[idx, 1], wp[idx, 2], wp[idx, 3] + np.pi / 2], axis=1)
    if lateral:
        off = np.random.default_rng(seed).uniform(-lateral, lateral, n)
        poses[:, 0] += -off * np.sin(poses[:, 2])
        poses[:, 1] += off * np.cos(poses[:, 2])
    return poses


def resampled_map(size, base_map=None, map_ext=".png"):
    """BASELINE config 5: example_map resampled with PIL NEAREST to size x size cells covering the same
    100 m x 100 m, re-thresholded like set_map, EDT by scipy.  Returns a MapData."""
    import yaml
    from PIL import Image
    from scipy.ndimage import distance_transform_edt
    base_map = base_map or os.path.join(MAPS_DIR, "example_map.yaml")
    img = Image.open(os.path.splitext(base_map)[0] + map_ext).transpose(Image.FLIP_TOP_BOTTOM)
    with open(base_map, "r") as stream:
        meta = yaml.safe_load(stream)
    extent = float(meta["resolution"]) * img.size[0]
    occ = np.asarray(img.resize((size, size), Image.NEAREST), dtype=np.float64)
    res = extent / size
    edt = res * distance_transform_edt(occ > 128.0)
    return MapData.from_arrays(edt, res, tuple(float(o) for o in meta["origin"]))


def free_space_poses(m, n, seed=7, clearance=0.3):
    """n poses uniform over the cells of MapData m that are more than `clearance` metres from an obstacle,
    heading ~ U(0, 2 pi) (config 5)."""
    rng = np.random.default_rng(seed)
    free = np.argwhere(m.edt > clearance)
    pick = free[rng.integers(0, free.shape[0], n)]
    gx = (pick[:, 1] + rng.uniform(0, 1, n)) * m.resolution
    gy = (pick[:, 0] + rng.uniform(0, 1, n)) * m.resolution
    x = m.origin[0] + gx * m.orig_c - gy * m.orig_s
    y = m.origin[1] + gx * m.orig_s + gy * m.orig_c
    return np.stack([x, y, rng.uniform(0, 2 * np.pi, n)], axis=1)


def gpu_edt(free, resolution, device=0):
    """get_dt (laser_models.py:40-53) on the device: free = (h, w) array/tensor, non-zero = free space.
    Returns a (h, w) float64 CUDA tensor equal to resolution * scipy.ndimage.distance_transform_edt(free)."""
    import ctypes
    import torch
    from . import _native as nat
    dev = torch.device("cuda", int(device))
    f = (torch.as_tensor(free, device=dev) != 0).to(torch.uint8).contiguous()
    h, w = f.shape
    out = torch.empty((h, w), dtype=torch.float64, device=dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    nat.check(nat.load().f110_edt_build(ctypes.c_void_p(f.data_ptr()), h, w, float(resolution), ctypes.c_void_p(out.data_ptr()),
                                        int(device), stream))
    return out


def bind_to_gpu_numa(device_index):
    """Pin this process to the CPUs NVML reports as local to the GPU (its NUMA node), so that the page-locked host
    buffers allocated afterwards sit next to the GPU's PCIe root.  For one process per GPU on a multi-socket box:
    with 8 ranks copying 4.6 GB of scans per step each, host buffers on the wrong socket cap the aggregate PCIe rate.
    Returns the CPU list, or None when NVML / the affinity call is unavailable (nothing changes then)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        idx = device_index
        if vis:
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if device_index < len(ids) and ids[device_index].isdigit():
                idx = int(ids[device_index])
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * i + b for i, w in enumerate(mask) for b in range(64) if (int(w) >> b) & 1]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None


class NoiseRowsAhead(object):
    """The parity-mode noise stream, drawn ahead of the steps by a helper thread (pure numpy: no CUDA call leaves the caller's
    thread).  Row k is the k-th `default_rng(seed).normal(0, 0.01, B)` of a car since its reset (base_classes.py:202,
    laser_models.py:451); a (count, B) draw fills row by row, so chunks of rows are the same numbers as successive B-draws.
    PCG64 + ziggurat is sequential -- 17 us per row of 1080 normals, a third of a single-env step when drawn in the stepping
    thread -- but numpy releases the GIL while it fills, so a second thread keeps two chunks ready.  take(n) returns the
    next n rows of the stream, in order, waiting for the helper only if the caller outruns it."""
    CHUNK = 128

    def __init__(self, seed, num_beams, std=0.01):
        import queue
        import threading
        self._rng = np.random.default_rng(seed=seed)
        self._beams, self._std = int(num_beams), float(std)
        self._q = queue.Queue(maxsize=2)
        self._stop = threading.Event()
        self._left = None
        self._threaded = os.environ.get("F110_NOISE_AHEAD", "1") != "0"      # 0: draw in the caller's thread, at take()
        if self._threaded:
            self._t = threading.Thread(target=self._run, name="f110-noise-rows", daemon=True)
            self._t.start()
            _live_noise_threads.add(self)

    def _run(self):
        import queue
        while not self._stop.is_set():
            rows = self._rng.normal(0., self._std, size=(self.CHUNK, self._beams))
            while not self._stop.is_set():
                try:
                    self._q.put(rows, timeout=0.2)
                    break
                except queue.Full:
                    pass

    def take(self, n):
        parts, have = [], 0
        if self._left is not None:
            parts.append(self._left); have = self._left.shape[0]; self._left = None
        while have < n:
            rows = self._q.get() if self._threaded else self._rng.normal(0., self._std, size=(max(n - have, 1), self._beams))
            parts.append(rows); have += rows.shape[0]
        out = parts[0] if len(parts) == 1 else np.concatenate(parts, axis=0)
        if have > n:
            self._left = out[n:]
            out = out[:n]
        return np.ascontiguousarray(out)

    def close(self, join=False):
        self._stop.set()
        try:
            while True:
                self._q.get_nowait()
        except Exception:
            pass
        if join and self._threaded:
            self._t.join(timeout=1.0)


# helper threads still alive at interpreter exit are stopped and joined first: a daemon thread that comes back from numpy
# for the GIL while the interpreter is being torn down is the one thing that could go wrong with them
import atexit      # noqa: E402
import weakref     # noqa: E402
_live_noise_threads = weakref.WeakSet()


@atexit.register
def _stop_noise_threads():
    for a in list(_live_noise_threads):
        a.close(join=True)
