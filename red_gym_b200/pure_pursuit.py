"""Batched pure-pursuit planner on the GPU: the reference's `PurePursuitPlanner` (examples/waypoint_follow.py:143-217)
for N cars at once.  `plan` has the reference's signature and return order (speed, steering_angle); `plan_actions`
writes the (N, A, 2) [steer, speed] action tensor `BatchF110Env.step` takes, straight from the env's device state.
Runs in libf110_b200.so (csrc/f110_pursuit.cu); no CPU path."""
import ctypes

import numpy as np

from . import _native as nat


class BatchPurePursuitPlanner(object):
    def __init__(self, conf, wb, device=0):
        """conf: the reference's config namespace (wpt_path, wpt_delim, wpt_rowskip, wpt_xind, wpt_yind, wpt_vind), or an
        (n, >=3) array whose columns are (x, y, speed); wb: wheelbase."""
        import torch
        self._torch = torch
        self.wheelbase = float(wb)
        self.conf = conf
        self.max_reacquire = 20.
        if hasattr(conf, "wpt_path"):
            self.waypoints = np.loadtxt(conf.wpt_path, delimiter=conf.wpt_delim, skiprows=conf.wpt_rowskip)   # load_waypoints (:160-164)
            cols = (conf.wpt_xind, conf.wpt_yind, conf.wpt_vind)
        else:
            self.waypoints = np.asarray(conf, dtype=np.float64)
            cols = (0, 1, 2)
        wx, wy, wv = (np.ascontiguousarray(self.waypoints[:, c], dtype=np.float64) for c in cols)
        self.device = torch.device("cuda", int(device))
        self.lib = nat.load()
        h = ctypes.c_void_p()
        p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
        nat.check(self.lib.f110_pursuit_create(p(wx), p(wy), p(wv), wx.shape[0], self.wheelbase, self.max_reacquire, int(device),
                                               ctypes.byref(h)))
        self._h = h

    def _run(self, x, y, theta, n, stride, lookahead_distance, vgain, out, picked=None):
        torch = self._torch
        ld_q = vg_q = None
        if isinstance(lookahead_distance, torch.Tensor):
            ld_q = lookahead_distance.to(self.device, torch.float64).contiguous(); lookahead_distance = 0.0
        if isinstance(vgain, torch.Tensor):
            vg_q = vgain.to(self.device, torch.float64).contiguous(); vgain = 0.0
        vp = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        nat.check(self.lib.f110_pursuit_plan(self._h, vp(x), vp(y), vp(theta), n, stride, float(lookahead_distance), float(vgain),
                                             vp(ld_q), vp(vg_q), vp(out), vp(picked), stream))
        return out

    def plan(self, pose_x, pose_y, pose_theta, lookahead_distance, vgain, return_picked=False):
        """PurePursuitPlanner.plan (:203-217) elementwise over tensors/arrays of poses -> (speed, steering_angle)."""
        torch = self._torch
        t = lambda v: torch.as_tensor(np.asarray(v) if not isinstance(v, torch.Tensor) else v, dtype=torch.float64,
                                      device=self.device).reshape(-1).contiguous()
        x, y, th = t(pose_x), t(pose_y), t(pose_theta)
        n = x.shape[0]
        out = torch.empty((n, 2), dtype=torch.float64, device=self.device)
        picked = torch.empty((n, 2), dtype=torch.int32, device=self.device) if return_picked else None
        self._run(x, y, th, n, 1, lookahead_distance, vgain, out, picked)
        if return_picked:
            return out[:, 1], out[:, 0], picked
        return out[:, 1], out[:, 0]

    def plan_actions(self, env, lookahead_distance, vgain, out=None):
        """Actions for every agent of a BatchF110Env from its device state: (N, A, 2) [steer, speed]."""
        torch = self._torch
        n = env.num_envs * env.num_agents
        if out is None:
            out = torch.empty((env.num_envs, env.num_agents, 2), dtype=torch.float64, device=self.device)
        s = env.state                      # (7, N, A): rows x, y, steer, v, yaw, ...
        self._run(s[0], s[1], s[4], n, 1, lookahead_distance, vgain, out)
        return out

    def close(self):
        h = getattr(self, "_h", None)
        if h:
            self.lib.f110_pursuit_destroy(h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
