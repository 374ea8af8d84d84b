"""ctypes front-end of the CPU oracle (oracle/f110_oracle.c).  TEST INFRASTRUCTURE, NOT PRODUCT.

Self-contained on purpose: map loading, trig/beam tables and the noise stream are restated
here (from the reference, cited below) rather than imported from the product package, so the
product is checked against something that shares none of its code.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libf110_oracle.so")

PARAM_KEYS = ["mu", "C_Sf", "C_Sr", "lf", "lr", "h", "m", "I", "s_min", "s_max", "sv_min", "sv_max",
              "v_switch", "a_max", "v_min", "v_max", "width", "length"]
# f110_env.py:128
DEFAULT_PARAMS = {"mu": 1.0489, "C_Sf": 4.718, "C_Sr": 5.4562, "lf": 0.15875, "lr": 0.17145, "h": 0.074,
                  "m": 3.74, "I": 0.04712, "s_min": -0.4189, "s_max": 0.4189, "sv_min": -3.2, "sv_max": 3.2,
                  "v_switch": 7.319, "a_max": 9.51, "v_min": -5.0, "v_max": 20.0, "width": 0.31, "length": 0.58}


def build(force=False):
    src = os.path.join(HERE, "f110_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", HERE, "-B"], check=True, capture_output=True)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(LIB_PATH)
        vp, i32, f64, i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_int64
        L.f110o_create.restype = vp
        L.f110o_create.argtypes = [i32] * 6 + [f64] * 5
        L.f110o_destroy.argtypes = [vp]
        L.f110o_set_map.argtypes = [vp, vp, i32, i32, f64, f64, f64, f64, f64]
        L.f110o_set_tables.argtypes = [vp] * 6
        L.f110o_set_params.argtypes = [vp, vp, i32]
        L.f110o_reset.argtypes = [vp, vp, vp, vp, i64, i32]
        L.f110o_step.argtypes = [vp, vp, vp, i64, i32]
        L.f110o_scan.argtypes = [vp, vp, i32, vp, vp, i32]
        L.f110o_gjk.argtypes = [vp, vp]
        L.f110o_gjk.restype = i32
        L.f110o_collision_multiple.argtypes = [i32, vp, vp, vp]
        L.f110o_get_vertices.argtypes = [vp, f64, f64, vp]
        L.f110o_ray_cast.argtypes = [vp, vp, vp, vp]
        L.f110o_check_ttc.argtypes = [vp, vp, f64]
        L.f110o_check_ttc.restype = i32
        L.f110o_dynamics_st.argtypes = [vp] * 4
        L.f110o_pid.argtypes = [f64] * 8 + [vp, vp]
        for name in ("state", "scans", "collisions", "collision_idx", "toggles", "lap_counts", "lap_times",
                     "cur_time", "done", "in_collision", "noise_step"):
            getattr(L, "f110o_" + name).restype = vp
            getattr(L, "f110o_" + name).argtypes = [vp]
        L.f110o_lookups.restype = ctypes.c_uint64
        L.f110o_lookups.argtypes = [vp, i32]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


# ---------------------------------------------------------------------------------- host setup
def load_map(map_path, map_ext, origin_yaw=None):
    """ScanSimulator2D.set_map (laser_models.py:383-427): flip, threshold at 128, yaml, scipy EDT.
    origin_yaw: use this instead of the yaml's origin[2] (every bundled yaml has 0; the rotated goldens of
    oracle/gen_golden.py were made from a yaml that differs in this one number)."""
    import yaml
    from PIL import Image
    from scipy.ndimage import distance_transform_edt
    img_path = os.path.splitext(map_path)[0] + map_ext
    img = np.array(Image.open(img_path).transpose(Image.FLIP_TOP_BOTTOM)).astype(np.float64)
    img[img <= 128.] = 0.
    img[img > 128.] = 255.
    with open(map_path, "r") as f:
        meta = yaml.safe_load(f)
    res = meta["resolution"]
    origin = list(meta["origin"])
    if origin_yaw is not None:
        origin[2] = float(origin_yaw)
    dt = res * distance_transform_edt(img)        # laser_models.py:40-53
    return dict(dt=np.ascontiguousarray(dt), h=img.shape[0], w=img.shape[1], res=float(res),
                ox=float(origin[0]), oy=float(origin[1]), oc=float(np.cos(origin[2])), os=float(np.sin(origin[2])))


def trig_tables(theta_dis=2000):
    """laser_models.py:379-381 -- linspace WITH endpoint."""
    th = np.linspace(0.0, 2 * np.pi, num=theta_dis)
    return np.sin(th), np.cos(th)


def beam_tables(num_beams, fov, params):
    """base_classes.py:120-156: scan_angles, cosines, side_distances."""
    inc = fov / (num_beams - 1)
    cosines = np.zeros((num_beams,))
    scan_angles = np.zeros((num_beams,))
    side = np.zeros((num_beams,))
    dist_sides = params["width"] / 2.
    dist_fr = (params["lf"] + params["lr"]) / 2.
    for i in range(num_beams):
        angle = -fov / 2. + i * inc
        scan_angles[i] = angle
        cosines[i] = np.cos(angle)
        if angle > 0:
            if angle < np.pi / 2:
                to_side = dist_sides / np.sin(angle); to_fr = dist_fr / np.cos(angle)
            else:
                to_side = dist_sides / np.cos(angle - np.pi / 2.); to_fr = dist_fr / np.sin(angle - np.pi / 2.)
        else:
            if angle > -np.pi / 2:
                to_side = dist_sides / np.sin(-angle); to_fr = dist_fr / np.cos(-angle)
            else:
                to_side = dist_sides / np.cos(-angle - np.pi / 2); to_fr = dist_fr / np.sin(-angle - np.pi / 2)
        side[i] = min(to_side, to_fr)
    return scan_angles, cosines, side


def noise_table(seed, rows, num_beams, std_dev=0.01):
    """Row k = the k-th rng.normal(0., std, size=B) draw after reset (base_classes.py:202, laser_models.py:451)."""
    rng = np.random.default_rng(seed=seed)
    return np.stack([rng.normal(0., std_dev, size=num_beams) for _ in range(rows)])


def params_vec(params):
    return np.array([params[k] for k in PARAM_KEYS], dtype=np.float64)


# ---------------------------------------------------------------------------------- batch env
class OracleEnv(object):
    """N independent copies of the reference's F110Env (f110_env.py:53) stepped on the CPU."""

    def __init__(self, num_envs=1, num_agents=2, map=None, map_ext=".png", params=None, timestep=0.01, ego_idx=0,
                 integrator="RK4", fov=2 * np.pi, seed=12345, num_beams=1080, noise=True, noise_rows=2048,
                 nthreads=0, map_data=None, origin_yaw=None):
        with np.errstate(divide="ignore"):
            self.params = dict(DEFAULT_PARAMS if params is None else params)
            self.N, self.A, self.B = num_envs, num_agents, num_beams
            self.nthreads = nthreads
            self.seed = seed
            L = lib()
            integ = {"RK4": 1, "Euler": 2}[integrator if isinstance(integrator, str) else integrator.name]
            self.h = L.f110o_create(num_envs, num_agents, num_beams, integ, ego_idx, 2000, timestep, fov, 1e-4, 30.0, 0.005)
            if not self.h:
                raise ValueError("f110o_create failed")
            self.map = map_data if map_data is not None else load_map(map + ".yaml" if not map.endswith(".yaml") else map, map_ext, origin_yaw)
            m = self.map
            L.f110o_set_map(self.h, _p(m["dt"]), m["h"], m["w"], m["res"], m["ox"], m["oy"], m["oc"], m["os"])
            s, c = trig_tables(2000)
            sa, bc, sd = beam_tables(num_beams, fov, self.params)
            self.scan_angles = sa
            self._tables = (s, c, sa, bc, sd)
            L.f110o_set_tables(self.h, _p(s), _p(c), _p(sa), _p(bc), _p(sd))
            pv = params_vec(self.params)
            for a in range(num_agents + 1):
                L.f110o_set_params(self.h, _p(pv), a)
            self.noise = noise_table(seed, noise_rows, num_beams) if noise else None
            self.noise_rows = noise_rows if noise else 0
        na = num_envs * num_agents

        def view(name, shape, dtype=np.float64):
            ptr = getattr(L, "f110o_" + name)(self.h)
            n = int(np.prod(shape))
            ct = {np.float64: ctypes.c_double, np.uint8: ctypes.c_uint8, np.int64: ctypes.c_int64}[dtype]
            return np.ctypeslib.as_array(ctypes.cast(ptr, ctypes.POINTER(ct)), shape=(n,)).reshape(shape)

        self.state = view("state", (num_envs, num_agents, 7))
        self.scans = view("scans", (num_envs, num_agents, num_beams))
        self.collisions = view("collisions", (num_envs, num_agents))
        self.collision_idx = view("collision_idx", (num_envs, num_agents))
        self.toggles = view("toggles", (num_envs, num_agents))
        self.lap_counts = view("lap_counts", (num_envs, num_agents))
        self.lap_times = view("lap_times", (num_envs, num_agents))
        self.cur_time = view("cur_time", (num_envs,))
        self.done = view("done", (num_envs,), np.uint8)
        self.in_collision = view("in_collision", (num_envs, num_agents), np.uint8)
        self.noise_step = view("noise_step", (num_envs,), np.int64)

    def set_params(self, params, agent_idx=-1):
        pv = params_vec(params)
        idxs = range(self.A) if agent_idx < 0 else [agent_idx]
        for a in idxs:
            lib().f110o_set_params(self.h, _p(pv), a)

    def _ensure_noise(self):
        if self.noise is not None and int(self.noise_step.max()) + 2 > self.noise_rows:
            self.noise_rows *= 2
            self.noise = noise_table(self.seed, self.noise_rows, self.B)

    def reset(self, poses, mask=None):
        poses = np.ascontiguousarray(np.asarray(poses, dtype=np.float64).reshape(self.N, self.A, 3))
        m = None if mask is None else np.ascontiguousarray(np.asarray(mask, dtype=np.uint8))
        lib().f110o_reset(self.h, _p(poses), _p(m), _p(self.noise), self.noise_rows, self.nthreads)

    def step(self, actions):
        self._ensure_noise()
        actions = np.ascontiguousarray(np.asarray(actions, dtype=np.float64).reshape(self.N, self.A, 2))
        lib().f110o_step(self.h, _p(actions), _p(self.noise), self.noise_rows, self.nthreads)

    def scan(self, poses, return_lookups=False):
        poses = np.ascontiguousarray(np.asarray(poses, dtype=np.float64).reshape(-1, 3))
        n = poses.shape[0]
        out = np.empty((n, self.B))
        lk = np.zeros((n,), dtype=np.uint64)
        lib().f110o_scan(self.h, _p(poses), n, _p(out), _p(lk), self.nthreads)
        return (out, lk) if return_lookups else out

    def lookups(self, clear=True):
        return int(lib().f110o_lookups(self.h, 1 if clear else 0))

    def close(self):
        if getattr(self, "h", None):
            lib().f110o_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
