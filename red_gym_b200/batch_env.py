"""BatchF110Env: N independent copies of the reference's F110Env stepped by one GPU.

Same semantics as `F110Env` (f110_env.py:53-416) per env -- same kwargs, same params dict, same
observation keys -- with a leading env dimension and torch CUDA tensors instead of python lists.
All arithmetic happens in libf110_b200.so (hand-written sm_100a kernels behind a C ABI); this
class only owns the handle, uploads setup data and wraps the library's device buffers as
zero-copy torch tensors.
"""
import ctypes

import numpy as np

from . import _native as nat
from . import host


def _integrator_code(integrator):
    name = integrator if isinstance(integrator, str) else getattr(integrator, "name", None)
    if name == "RK4":
        return nat.INTEGRATOR_RK4
    if name == "Euler":
        return nat.INTEGRATOR_EULER
    # base_classes.py:395-396
    raise SyntaxError("Invalid Integrator Specified. Provided %s. Please choose RK4 or Euler" % (name,))


def _hp(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


class BatchF110Env(object):
    """
    Args (reference kwargs, f110_env.py:100-157): seed, map, map_ext, params, num_agents, timestep,
        ego_idx, integrator, fov.
    Batch kwargs: num_envs; device (CUDA ordinal); dtype 'fp64' (parity path) | 'fp32' (fast lidar);
        noise 'numpy' (host PCG64 rows, parity) | 'philox' (device) | 'off'; num_beams (1080);
        auto_reset + reset_poses (pool [P, A, 3]) for throughput runs; env_offset = global id of env 0 (sharding);
        origin_yaw: replaces origin[2] of the map yaml (map-frame rotation, laser_models.py:421-422);
        scan_format 'native' | 'u16' (fp32 path, one agent: scans stored as uint16 at 0.5 mm per LSB, see F110_SCAN_U16).

    Aliasing contract: every tensor this class returns (`scans`, `state`, the observation dict of reset/step, `done`, ...)
    is a zero-copy VIEW of a device buffer the library owns and overwrites in place on the next reset/step; keep a
    `.clone()` of whatever must outlive the step (replay buffers, n-step returns), or pass `copy=True` to step/reset.
    After close() the views are dropped (the memory behind them is freed).
    """

    def __init__(self, num_envs=1, device=0, dtype="fp64", noise=None, num_beams=1080, auto_reset=False,
                 reset_poses=None, raycast_variant=0, map_data=None, env_offset=0, origin_yaw=None, scan_format="native", **kwargs):
        import torch
        self._torch = torch
        self.lib = nat.load()
        self.seed = kwargs.get("seed", 12345)
        self.map_name = kwargs.get("map", None)
        self.map_path = host.resolve_map(self.map_name)
        self.map_ext = kwargs.get("map_ext", ".png")
        self.params = dict(kwargs.get("params", host.DEFAULT_PARAMS))
        self.num_agents = kwargs.get("num_agents", 2)
        self.timestep = kwargs.get("timestep", 0.01)
        self.ego_idx = kwargs.get("ego_idx", 0)
        self.integrator = kwargs.get("integrator", "RK4")
        self.fov = kwargs.get("fov", 2 * np.pi)
        # ScanSimulator2D's own constructor arguments (laser_models.py:360); the env always uses the defaults
        self.eps = float(kwargs.get("eps", 1e-4))
        self.theta_dis = int(kwargs.get("theta_dis", 2000))
        self.max_range = float(kwargs.get("max_range", 30.0))
        self.num_envs = int(num_envs)
        self.num_beams = int(num_beams)
        self.device_index = int(device)
        self.dtype = dtype
        if dtype not in ("fp64", "fp32"):
            raise ValueError("dtype must be 'fp64' or 'fp32'")
        if noise is None:
            noise = "numpy" if dtype == "fp64" else "philox"
        noise_mode = {"off": nat.NOISE_OFF, "numpy": nat.NOISE_TABLE, "philox": nat.NOISE_PHILOX}[noise]
        self.noise = noise
        if scan_format not in ("native", "u16"):
            raise ValueError("scan_format must be 'native' or 'u16'")
        self.scan_format = scan_format
        cfg = nat.Config(abi_version=nat.F110_ABI_VERSION, n_envs=self.num_envs, n_agents=self.num_agents,
                         n_beams=self.num_beams, integrator=_integrator_code(self.integrator), ego_idx=self.ego_idx,
                         theta_dis=self.theta_dis, precision=nat.PRECISION_FP64 if dtype == "fp64" else nat.PRECISION_FP32,
                         noise_mode=noise_mode, auto_reset=1 if auto_reset else 0, raycast_variant=raycast_variant, env_offset=int(env_offset),
                         scan_format=nat.SCAN_U16 if scan_format == "u16" else nat.SCAN_NATIVE, reserved0=0,
                         timestep=self.timestep, fov=self.fov, eps=self.eps, max_range=self.max_range, ttc_thresh=0.005,
                         noise_std=0.01, seed=self.seed)
        h = ctypes.c_void_p()
        nat.check(self.lib.f110_create(ctypes.byref(cfg), self.device_index, ctypes.byref(h)))
        self._h = h
        self.device = torch.device("cuda", self.device_index)
        self._noise_rows = 0
        self._noise_rng = None
        self._steps_since_full_reset = 0
        # setup uploads
        sines, cosines = host.trig_tables(self.theta_dis)
        self.scan_angles, beam_cos, side = host.beam_tables(self.num_beams, self.fov, self.params)
        nat.check(self.lib.f110_set_tables(h, _hp(sines), _hp(cosines), _hp(self.scan_angles), _hp(beam_cos), _hp(side), None))
        pv = host.params_vector(self.params)
        nat.check(self.lib.f110_set_params(h, _hp(pv), -1, None))
        nat.check(self.lib.f110_set_params(h, _hp(pv), self.num_agents, None))
        self.map = None
        if map_data is not None:
            self._upload_map(map_data)
        else:
            self._upload_map(host.MapData(self.map_path, self.map_ext, edt=kwargs.get("edt", "gpu"), origin_yaw=origin_yaw))
        if self.noise == "numpy":
            self._ensure_noise(64)
        if reset_poses is not None:
            self.set_reset_pool(reset_poses)
        self._wrap_buffers()

    # ------------------------------------------------------------------ setup
    def _stream(self):
        # torch's current stream of this device as a raw cudaStream_t.  torch.cuda.current_stream() builds a Stream object
        # per call (~8 us, a tenth of a single-env step); the raw accessor behind it is used where this torch has it.
        raw = getattr(self._torch._C, "_cuda_getCurrentRawStream", None)
        if raw is not None:
            return ctypes.c_void_p(raw(self.device_index))
        return ctypes.c_void_p(self._torch.cuda.current_stream(self.device).cuda_stream)

    def _upload_map(self, m):
        if m.edt is None:      # thresholded image only: get_dt runs on the device (csrc/f110_edt.cu)
            nat.check(self.lib.f110_set_map_image(self._h, _hp(m.free), m.height, m.width, m.resolution, m.origin[0], m.origin[1],
                                                  m.orig_c, m.orig_s, None))
        else:
            nat.check(self.lib.f110_set_map(self._h, _hp(m.edt), m.height, m.width, m.resolution, m.origin[0], m.origin[1],
                                            m.orig_c, m.orig_s, None))
        self.map = m          # only after the library accepted it: a rejected update leaves the previous map in place

    def _ensure_noise(self, rows, exact=False):
        """Parity mode: the device table must hold the first `rows` draws of the per-car noise stream (row k = the k-th
        `default_rng(seed).normal(0, 0.01, B)` after a reset, base_classes.py:202, laser_models.py:451).  PCG64 + ziggurat is
        a sequential stream: the generator is kept and only the missing rows are drawn (a (count, B) draw fills row by row,
        the same numbers as count successive B-draws) and uploaded -- one row per step amortised, as in the reference.
        Lazy growth (from step) comes in chunks of 128 rows, 2 ms of numpy each, uploaded beside the step's stream: the steps
        already queued cover it, where a table that doubled stalled the device for as long as the draw took (75 ms at 4096 rows)."""
        if self.noise != "numpy" or rows <= self._noise_rows:
            return
        target = int(rows) if exact else max(int(rows), self._noise_rows + 128)
        if self._noise_rng is None:
            self._noise_rng = host.NoiseRowsAhead(self.seed, self.num_beams)       # the rows are drawn ahead by a helper thread
        new = self._noise_rng.take(target - self._noise_rows)
        nat.check(self.lib.f110_set_noise_rows(self._h, _hp(new), self._noise_rows, new.shape[0], self._stream()))
        self._noise_rows = target

    def _wrap_buffers(self):
        torch = self._torch
        b = nat.Buffers()
        nat.check(self.lib.f110_get_buffers(self._h, ctypes.byref(b)))
        N, A, B = self.num_envs, self.num_agents, self.num_beams

        def t(ptr, shape, typestr):
            return torch.as_tensor(nat.DevArray(ptr, shape, typestr, self), device=self.device)

        self.state = t(b.state, (7, N, A), "<f8")
        self.obs_block = t(b.obs_block, (10, N, A), "<f8")
        self.collisions = t(b.collisions, (N, A), "<f8")
        self.collision_idx = t(b.collision_idx, (N, A), "<f8")
        self.lap_times = t(b.lap_times, (N, A), "<f8")
        self.lap_counts = t(b.lap_counts, (N, A), "<f8")
        self.toggle_list = t(b.toggles, (N, A), "<i4")
        self.done = t(b.done, (N,), "|u1")
        self.in_collision = t(b.in_collision, (N, A), "|u1")
        self.scans = t(b.scans, (N, A, B), "<u2" if self.scan_format == "u16" else "<f8" if self.dtype == "fp64" else "<f4")
        self.current_time = t(b.cur_time, (N,), "<f8")
        self.noise_step = t(b.noise_step, (N,), "<i8")
        self.actions = t(b.actions, (N, A, 2), "<f8")
        self.episode_count = t(b.episode_count, (N,), "<i8")
        self._lookups = t(b.lookups, (1,), "<i8")
        # bool views of the byte flags: step() then launches no torch kernel at all
        self.checkpoint_done = t(b.checkpoint_done, (N, A), "|u1").view(torch.bool)
        self.done_bool = self.done.view(torch.bool)
        self._zeros_na = torch.zeros((N, A), dtype=torch.float64, device=self.device)
        st = self.state
        self._obs_proto = {"ego_idx": self.ego_idx, "scans": self.scans, "poses_x": st[0], "poses_y": st[1], "poses_theta": st[4],
                           "linear_vels_x": st[3], "linear_vels_y": self._zeros_na, "ang_vels_z": st[5],
                           "collisions": self.collisions, "lap_times": self.lap_times, "lap_counts": self.lap_counts}
        self._n_act = N * A * 2
        self._noise_is_numpy = self.noise == "numpy"
        self._c_step = self.lib.f110_step
        self._c_step_host = self.lib.f110_step_host
        raw = getattr(torch._C, "_cuda_getCurrentRawStream", None)
        self._raw_stream = raw if raw is not None else (lambda idx: torch.cuda.current_stream(self.device).cuda_stream)

    def set_reset_pool(self, poses):
        poses = np.ascontiguousarray(np.asarray(poses, dtype=np.float64).reshape(-1, self.num_agents, 3))
        nat.check(self.lib.f110_set_reset_pool(self._h, _hp(poses), poses.shape[0], None))

    # ------------------------------------------------------------------ reference API (batched)
    def update_map(self, map_path, map_ext, edt="gpu"):
        """F110Env.update_map (f110_env.py:349-360): map_path is the yaml path.  The EDT (get_dt) is built on the device by
        default (bit-identical to scipy's, milliseconds instead of seconds); edt='scipy' does it on the host."""
        self._upload_map(host.MapData(map_path, map_ext, edt=edt))

    def update_params(self, params, index=-1):
        """F110Env.update_params -> Simulator.update_params (f110_env.py:362, base_classes.py:507-527)."""
        if index >= self.num_agents:
            raise IndexError("Index given is out of bounds for list of agents.")
        pv = host.params_vector(params)
        nat.check(self.lib.f110_set_params(self._h, _hp(pv), -1 if index < 0 else index, None), IndexError)

    def reset(self, poses, mask=None, copy=False, zero_action_step=True):
        """F110Env.reset (f110_env.py:304-347).  poses: (N, A, 3) [x, y, yaw] host array/tensor.
        mask: optional (N,) bool -- only those envs are reset (the others are untouched).
        zero_action_step=False: Simulator.reset only (base_classes.py:607-623) -- poses, FIFOs, flags and the noise stream
        are reset, the zero-action step F110Env.reset adds (:335-336) is left to the caller.
        Returns views of library buffers (see the class docstring); copy=True returns clones."""
        poses = np.asarray(poses.cpu() if hasattr(poses, "cpu") else poses, dtype=np.float64)
        if poses.ndim == 2 and self.num_envs == 1:
            poses = poses[None]
        if poses.shape[0] != self.num_envs:
            raise ValueError("Number of pose sets for reset does not match number of envs.")
        if poses.shape[1] != self.num_agents:
            raise ValueError("Number of poses for reset does not match number of agents.")   # base_classes.py:618-619
        poses = np.ascontiguousarray(poses)
        rot = host.start_rotation(poses[:, self.ego_idx, 2])
        m = None
        if mask is not None:
            m = np.ascontiguousarray(np.asarray(mask.cpu() if hasattr(mask, "cpu") else mask).astype(np.uint8))
        else:
            self._steps_since_full_reset = 0
        self._ensure_noise(2)
        if not zero_action_step:
            nat.check(self.lib.f110_reset_poses(self._h, _hp(poses), _hp(rot), _hp(m), self._stream()))
            return None
        nat.check(self.lib.f110_reset(self._h, _hp(poses), _hp(rot), _hp(m), self._stream()))
        self._steps_since_full_reset += 1
        return self._obs(copy)

    def step(self, action, copy=False):
        """F110Env.step (f110_env.py:261-302).  action: (N, A, 2) [steer, speed]; a CUDA float64 tensor is used
        in place, anything else is copied to the device first.
        The observation holds VIEWS of library buffers that the next step overwrites (class docstring);
        copy=True returns clones instead (one extra device copy of the scans per step)."""
        torch = self._torch
        # the common case first and cheaply: at a few thousand envs the step is ~0.1 ms of GPU work and every
        # microsecond of Python in this function shows in the throughput
        if type(action) is torch.Tensor and action.is_cuda and action.dtype is torch.float64 and action.is_contiguous():
            if action.numel() != self._n_act:
                raise ValueError("action must have shape (num_envs, num_agents, 2)")
            ptr = action.data_ptr()
        else:
            if not isinstance(action, torch.Tensor):
                action = torch.from_numpy(np.ascontiguousarray(np.asarray(action, dtype=np.float64)))
            if action.numel() != self._n_act:
                raise ValueError("action must have shape (num_envs, num_agents, 2)")
            self.actions.copy_(action.reshape(self.num_envs, self.num_agents, 2), non_blocking=True)
            ptr = None
        self._steps_since_full_reset += 1
        if self._noise_is_numpy and self._steps_since_full_reset + 1 > self._noise_rows:
            self._ensure_noise(self._steps_since_full_reset + 1)
        rc = self._c_step(self._h, ptr, self._raw_stream(self.device_index))
        if rc:
            nat.check(rc)
        return self._obs(copy)

    def _obs(self, copy=False):
        if copy:
            obs, reward, done, info = self._obs(False)
            obs = {k: (v.clone() if hasattr(v, "clone") else v) for k, v in obs.items()}
            return obs, reward, done.clone(), {"checkpoint_done": info["checkpoint_done"].clone()}
        # the views never change (fixed library buffers): built once in _wrap_buffers, handed out as fresh dicts
        return dict(self._obs_proto), self.timestep, self.done_bool, {"checkpoint_done": self.checkpoint_done}

    def step_host(self, actions, obs_block_out, scans_out=None, done_out=None):
        """One step through f110_step_host: host actions in, host observation out (pinned numpy/torch buffers)."""
        def ptr(x):
            if x is None:
                return None
            return x.data_ptr() if hasattr(x, "data_ptr") else x.ctypes.data
        self.step_host_ptrs(ptr(actions), ptr(obs_block_out), ptr(scans_out), ptr(done_out))

    def step_host_ptrs(self, actions_ptr, obs_ptr, scans_ptr=None, done_ptr=None):
        """step_host with the raw host addresses (ints) of buffers the caller keeps alive: for callers that reuse the same
        buffers every step (the drop-in Simulator) and want no per-call pointer work."""
        self._steps_since_full_reset += 1
        if self._noise_is_numpy and self._steps_since_full_reset + 1 > self._noise_rows:
            self._ensure_noise(self._steps_since_full_reset + 1)
        rc = self._c_step_host(self._h, actions_ptr, obs_ptr, scans_ptr, done_ptr, self._raw_stream(self.device_index))
        if rc:
            nat.check(rc)

    def render(self, mode="human", envs=None, width=1000, height=800, follow=True, camera=None):
        """F110Env.render (f110_env.py:385-416).  The reference's window modes 'human' / 'human_fast' need pyglet/OpenGL,
        which is unavailable on this box: they stay accepted no-ops.  mode='rgb_array' rasterises the same scene
        (rendering.py: map points, car quads, colours, orthographic camera) on the device and returns a
        (n, height, width, 3) uint8 CUDA tensor for the first `envs` environments (default: all).
        follow=True: camera on the ego car like examples/waypoint_follow.py:259-269; else `camera` =
        (left, right, bottom, top) in 50 x world units, default the reference's start-up window (zoom 1.2)."""
        assert mode in ["human", "human_fast", "rgb_array"]
        if mode != "rgb_array":
            return None
        torch = self._torch
        black = getattr(self.map, "black", None)
        if black is None:
            raise ValueError("render needs a map loaded from an image (map pixels == 0 are drawn)")
        if camera is None:
            camera = (-1.2 * width / 2, 1.2 * width / 2, -1.2 * height / 2, 1.2 * height / 2)
        key = (id(self.map), int(width), int(height), bool(follow), tuple(float(c) for c in camera))
        if getattr(self, "_render_key", None) != key:
            if getattr(self, "_render_h", None):
                self.lib.f110_render_destroy(self._render_h)
            cfg = nat.RenderConfig(abi_version=nat.F110_ABI_VERSION, width=int(width), height=int(height), n_agents=self.num_agents,
                                   ego_idx=self.ego_idx, map_h=self.map.height, map_w=self.map.width, resolution=self.map.resolution,
                                   origin_x=self.map.origin[0], origin_y=self.map.origin[1], follow_margin=800.0 if follow else -1.0,
                                   left=camera[0], right=camera[1], bottom=camera[2], top=camera[3])
            h = ctypes.c_void_p()
            nat.check(self.lib.f110_render_create(ctypes.byref(cfg), _hp(black), self.device_index, ctypes.byref(h)))
            self._render_h, self._render_key = h, key
        n = self.num_envs if envs is None else int(envs)
        out = torch.empty((n, int(height), int(width), 3), dtype=torch.uint8, device=self.device)
        s = self.state
        nat.check(self.lib.f110_render_frame(self._render_h, ctypes.c_void_p(s[0].data_ptr()), ctypes.c_void_p(s[1].data_ptr()),
                                             ctypes.c_void_p(s[4].data_ptr()), n, ctypes.c_void_p(out.data_ptr()), self._stream()))
        return out

    # ------------------------------------------------------------------ CUDA graphs
    def reserve_noise(self, steps):
        """Parity mode (noise='numpy'): upload the noise rows of the next `steps` steps now.  step() does this lazily from
        Python; a captured graph replays without Python, so the table has to cover every replayed step beforehand."""
        self._ensure_noise(self._steps_since_full_reset + int(steps) + 1, exact=True)

    def capture_step_graph(self, actions, steps=1, warmup=3):
        """One or more steps as a CUDA graph: f110_step enqueues only kernel launches on the caller's stream (no allocation, no
        synchronisation), so a small-batch step -- tens of microseconds of kernels inside as much Python and launch overhead
        per call -- can be captured once and replayed with a single launch per replay.

        actions: a CUDA float64 tensor (N, A, 2) -- every captured step reads it, the caller overwrites it in place between
        replays (e.g. from a policy that is part of a larger captured graph) -- or (steps, N, A, 2): captured step k reads
        actions[k].  Returns a torch.cuda.CUDAGraph; graph.replay() advances the env by `steps` steps, after which the usual views
        (self.scans, self.state, self.done, ...) hold the new observation.  `warmup` eager steps are taken first on a side stream
        (they DO advance the env).  Parity mode: call reserve_noise(total steps to be replayed) first; the episode step counter
        of this Python object is advanced by replayed(steps) only if the caller says so (note_replayed)."""
        torch = self._torch
        if not (isinstance(actions, torch.Tensor) and actions.is_cuda and actions.dtype == torch.float64 and actions.is_contiguous()):
            raise ValueError("actions must be a contiguous CUDA float64 tensor")
        per = self._n_act
        if actions.numel() == per:
            acts = [actions] * int(steps)
        elif actions.numel() == per * int(steps):
            acts = list(actions.reshape(int(steps), self.num_envs, self.num_agents, 2))
        else:
            raise ValueError("actions must have shape (num_envs, num_agents, 2) or (steps, num_envs, num_agents, 2)")
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for k in range(int(warmup)):
                self.step(acts[k % len(acts)])
        torch.cuda.current_stream(self.device).wait_stream(side)
        self.reserve_noise(len(acts))
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for a in acts:
                rc = self._c_step(self._h, a.data_ptr(), self._raw_stream(self.device_index))
                if rc:
                    nat.check(rc)
        return g

    def note_replayed(self, steps):
        """Tell the Python side that `steps` steps were taken by graph replays (keeps the lazy noise-row upload in step())."""
        self._steps_since_full_reset += int(steps)

    # ------------------------------------------------------------------ extras
    def raycast(self, poses, out=None):
        """ScanSimulator2D.scan(pose, None) for a batch of poses (n, 3) -> (n, num_beams)."""
        torch = self._torch
        poses = torch.as_tensor(poses, dtype=torch.float64, device=self.device).contiguous()
        n = poses.shape[0]
        if out is None:
            # (always float ranges, also for scan_format='u16' handles)
            out = torch.empty((n, self.num_beams), dtype=torch.float64 if self.dtype == "fp64" else torch.float32, device=self.device)
        nat.check(self.lib.f110_raycast(self._h, ctypes.c_void_p(poses.data_ptr()), n, ctypes.c_void_p(out.data_ptr()), self._stream()))
        return out

    def count_lookups(self, enable=True):
        nat.check(self.lib.f110_count_lookups(self._h, 1 if enable else 0))
        self._lookups.zero_()

    def lookups(self, clear=True):
        v = int(self._lookups.item())
        if clear:
            self._lookups.zero_()
        return v

    def enable_timing(self, enable=True):
        nat.check(self.lib.f110_enable_timing(self._h, 1 if enable else 0))

    def last_raycast_ms(self):
        ms = ctypes.c_float()
        nat.check(self.lib.f110_last_raycast_ms(self._h, ctypes.byref(ms)))
        return float(ms.value)

    @property
    def launch_count(self):
        return int(self.lib.f110_launch_count(self._h))

    def synchronize(self):
        self._torch.cuda.synchronize(self.device)

    def close(self):
        ahead = getattr(self, "_noise_rng", None)
        if ahead is not None:
            ahead.close()
            self._noise_rng = None
        rh = getattr(self, "_render_h", None)
        if rh:
            self.lib.f110_render_destroy(rh)
            self._render_h = None
        h = getattr(self, "_h", None)
        if h:
            # the tensors are views of memory f110_destroy frees: drop them (and the DevArray -> env reference cycle)
            for name in ("state", "obs_block", "collisions", "collision_idx", "lap_times", "lap_counts", "toggle_list", "done",
                         "in_collision", "scans", "current_time", "noise_step", "actions", "episode_count", "_lookups",
                         "checkpoint_done", "done_bool", "_obs_proto", "_zeros_na"):
                self.__dict__.pop(name, None)
            self.lib.f110_destroy(h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
