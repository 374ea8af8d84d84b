"""ctypes binding of libf110_b200.so (include/f110_b200.h).

There is no CPU fallback: if the library is missing, importing this module raises; if there is
no CUDA device, `f110_create` fails and the wrapper raises RuntimeError.
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# F110_B200_LIB selects another build of the same ABI (the bounds-checked libf110_b200_checked.so of tests/test_gpu_checked_build.py)
LIB_PATH = os.environ.get("F110_B200_LIB") or os.path.join(HERE, "libf110_b200.so")

F110_ABI_VERSION = 2
NUM_PARAMS = 18
PARAM_KEYS = ["mu", "C_Sf", "C_Sr", "lf", "lr", "h", "m", "I", "s_min", "s_max", "sv_min", "sv_max",
              "v_switch", "a_max", "v_min", "v_max", "width", "length"]
OK, ERR_INVALID, ERR_CUDA, ERR_STATE, ERR_NOMEM = 0, -1, -2, -3, -4
INTEGRATOR_RK4, INTEGRATOR_EULER = 1, 2
PRECISION_FP64, PRECISION_FP32 = 0, 1
NOISE_OFF, NOISE_TABLE, NOISE_PHILOX = 0, 1, 2
SCAN_NATIVE, SCAN_U16 = 0, 1
SCAN_U16_PER_METRE = 2000.0

# every symbol include/f110_b200.h declares (tests check the library exports all of them)
EXPORTS = ["f110_last_error", "f110_abi_version", "f110_device_count", "f110_create", "f110_destroy",
           "f110_set_map", "f110_set_map_image", "f110_edt_build", "f110_edt_build_noalloc", "f110_set_tables", "f110_set_params", "f110_set_noise_rows", "f110_set_reset_pool",
           "f110_reset", "f110_reset_poses", "f110_step", "f110_step_host", "f110_raycast", "f110_get_buffers", "f110_get_config",
           "f110_count_lookups", "f110_enable_timing", "f110_last_raycast_ms", "f110_l2_probe", "f110_build_is_checked", "f110_debug_violations", "f110_launch_count", "f110_synchronize",
           "f110_bitmap_create", "f110_bitmap_destroy", "f110_bitmap_render",
           "f110_pursuit_create", "f110_pursuit_destroy", "f110_pursuit_plan",
           "f110_render_create", "f110_render_destroy", "f110_render_frame",
           "f110_model_dynamics_st", "f110_model_dynamics_ks", "f110_model_pid", "f110_model_accl_constraints",
           "f110_model_steering_constraint", "f110_model_vertices", "f110_model_collision"]
BITMAP_MODES = {"FILL": 0, "POLYGON": 1, "RAYS": 2}


class Config(ctypes.Structure):
    _fields_ = [("abi_version", ctypes.c_int32), ("n_envs", ctypes.c_int32), ("n_agents", ctypes.c_int32),
                ("n_beams", ctypes.c_int32), ("integrator", ctypes.c_int32), ("ego_idx", ctypes.c_int32),
                ("theta_dis", ctypes.c_int32), ("precision", ctypes.c_int32), ("noise_mode", ctypes.c_int32),
                ("auto_reset", ctypes.c_int32), ("raycast_variant", ctypes.c_int32), ("env_offset", ctypes.c_int32),
                ("scan_format", ctypes.c_int32), ("reserved0", ctypes.c_int32),
                ("timestep", ctypes.c_double), ("fov", ctypes.c_double), ("eps", ctypes.c_double),
                ("max_range", ctypes.c_double), ("ttc_thresh", ctypes.c_double), ("noise_std", ctypes.c_double),
                ("seed", ctypes.c_uint64)]


class BitmapConfig(ctypes.Structure):
    _fields_ = [("abi_version", ctypes.c_int32), ("n_beams", ctypes.c_int32), ("target_beam_count", ctypes.c_int32),
                ("height", ctypes.c_int32), ("width", ctypes.c_int32), ("draw_mode", ctypes.c_int32),
                ("draw_center", ctypes.c_int32), ("bg_white", ctypes.c_int32), ("scan_is_f32", ctypes.c_int32),
                ("scaling_factor", ctypes.c_double)]


class RenderConfig(ctypes.Structure):
    _fields_ = [("abi_version", ctypes.c_int32), ("width", ctypes.c_int32), ("height", ctypes.c_int32), ("n_agents", ctypes.c_int32),
                ("ego_idx", ctypes.c_int32), ("map_h", ctypes.c_int32), ("map_w", ctypes.c_int32),
                ("resolution", ctypes.c_double), ("origin_x", ctypes.c_double), ("origin_y", ctypes.c_double),
                ("follow_margin", ctypes.c_double), ("left", ctypes.c_double), ("right", ctypes.c_double),
                ("bottom", ctypes.c_double), ("top", ctypes.c_double)]


class Buffers(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in
                ("state", "obs_block", "collisions", "collision_idx", "lap_times", "lap_counts", "toggles", "done",
                 "in_collision", "scans", "cur_time", "noise_step", "actions", "episode_count", "lookups", "checkpoint_done")]


_lib = None


def load():
    """Load the shared library (raises ImportError with a build hint when it is not there)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "red_gym_b200: %s is missing -- build it with `python red_gym_b200/build.py` "
            "(needs nvcc, targets sm_100a). There is no CPU fallback." % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    vp, i32, i64, f64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_double
    L.f110_last_error.restype = ctypes.c_char_p
    L.f110_last_error.argtypes = []
    L.f110_abi_version.restype = ctypes.c_int
    L.f110_device_count.restype = ctypes.c_int
    L.f110_create.argtypes = [ctypes.POINTER(Config), ctypes.c_int, ctypes.POINTER(vp)]
    L.f110_destroy.argtypes = [vp]
    L.f110_set_map.argtypes = [vp, vp, i32, i32, f64, f64, f64, f64, f64, vp]
    L.f110_set_map_image.argtypes = [vp, vp, i32, i32, f64, f64, f64, f64, f64, vp]
    L.f110_edt_build.argtypes = [vp, i32, i32, f64, vp, ctypes.c_int, vp]
    L.f110_edt_build_noalloc.argtypes = [vp, i32, i32, f64, vp, ctypes.c_int, vp]
    L.f110_set_tables.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    L.f110_set_params.argtypes = [vp, vp, i32, vp]
    L.f110_set_noise_rows.argtypes = [vp, vp, i64, i64, vp]
    L.f110_set_reset_pool.argtypes = [vp, vp, i32, vp]
    L.f110_reset.argtypes = [vp, vp, vp, vp, vp]
    L.f110_reset_poses.argtypes = [vp, vp, vp, vp, vp]
    L.f110_step.argtypes = [vp, vp, vp]
    L.f110_step_host.argtypes = [vp, vp, vp, vp, vp, vp]
    L.f110_raycast.argtypes = [vp, vp, i32, vp, vp]
    L.f110_get_buffers.argtypes = [vp, ctypes.POINTER(Buffers)]
    L.f110_get_config.argtypes = [vp, ctypes.POINTER(Config)]
    L.f110_count_lookups.argtypes = [vp, i32]
    L.f110_enable_timing.argtypes = [vp, i32]
    L.f110_last_raycast_ms.argtypes = [vp, ctypes.POINTER(ctypes.c_float)]
    L.f110_l2_probe.argtypes = [ctypes.c_int, i64, i32, vp, vp]
    L.f110_build_is_checked.argtypes = []
    L.f110_debug_violations.argtypes = [ctypes.c_int, vp, i32]
    L.f110_launch_count.argtypes = [vp]
    L.f110_launch_count.restype = i64
    L.f110_synchronize.argtypes = [vp, vp]
    L.f110_bitmap_create.argtypes = [ctypes.POINTER(BitmapConfig), vp, vp, vp, ctypes.c_int, ctypes.POINTER(vp)]
    L.f110_bitmap_destroy.argtypes = [vp]
    L.f110_bitmap_render.argtypes = [vp, vp, i64, vp, i32, vp]
    L.f110_pursuit_create.argtypes = [vp, vp, vp, i32, f64, f64, ctypes.c_int, ctypes.POINTER(vp)]
    L.f110_pursuit_destroy.argtypes = [vp]
    L.f110_pursuit_plan.argtypes = [vp, vp, vp, vp, i64, i64, f64, f64, vp, vp, vp, vp, vp]
    L.f110_render_create.argtypes = [ctypes.POINTER(RenderConfig), vp, ctypes.c_int, ctypes.POINTER(vp)]
    L.f110_render_destroy.argtypes = [vp]
    L.f110_render_frame.argtypes = [vp, vp, vp, vp, i64, vp, vp]
    L.f110_model_dynamics_st.argtypes = [vp, vp, vp, i64, vp, ctypes.c_int, vp]
    L.f110_model_dynamics_ks.argtypes = [vp, vp, vp, i64, vp, ctypes.c_int, vp]
    L.f110_model_pid.argtypes = [vp, f64, f64, f64, f64, i64, vp, ctypes.c_int, vp]
    L.f110_model_accl_constraints.argtypes = [vp, f64, f64, f64, f64, i64, vp, ctypes.c_int, vp]
    L.f110_model_steering_constraint.argtypes = [vp, f64, f64, f64, f64, i64, vp, ctypes.c_int, vp]
    L.f110_model_vertices.argtypes = [vp, f64, f64, i64, vp, ctypes.c_int, vp]
    L.f110_model_collision.argtypes = [vp, vp, i64, vp, ctypes.c_int, vp]
    for name in EXPORTS:
        fn = getattr(L, name)
        if name not in ("f110_last_error", "f110_launch_count"):
            fn.restype = ctypes.c_int
    if L.f110_abi_version() != F110_ABI_VERSION:
        raise ImportError("libf110_b200.so ABI %d != binding %d" % (L.f110_abi_version(), F110_ABI_VERSION))
    _lib = L
    return L


def last_error():
    return load().f110_last_error().decode("utf-8", "replace")


def check(rc, invalid_exc=ValueError):
    """Map a status code to the exception the reference raises for the same condition."""
    if rc == OK:
        return
    msg = last_error()
    if rc == ERR_INVALID:
        raise invalid_exc(msg)
    if rc == ERR_STATE:
        raise ValueError(msg)          # e.g. 'Map is not set for scan simulator.' (laser_models.py:445-446)
    if rc == ERR_NOMEM:
        raise MemoryError(msg)
    raise RuntimeError("libf110_b200: " + msg)


class DevArray(object):
    """Zero-copy view of library-owned device memory for torch.as_tensor (CUDA array interface v2)."""

    def __init__(self, ptr, shape, typestr, owner):
        self.owner = owner   # keeps the handle alive
        self.__cuda_array_interface__ = {"shape": tuple(int(s) for s in shape), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 2, "strides": None}
