"""The model functions of the step on their own, batched on the device: single-track / kinematic right-hand sides, the
speed/steer PID, car rectangles and GJK overlap (reference: gym/f110_gym/envs/dynamic_models.py:91-221,
collision_models.py:114-260).  They run the same device code as the step kernels (csrc/f110_kernels_fp64.cu), so the
reference's unit-test vectors (f_st_gt, the GJK known answers) can be checked on the GPU path directly.

Inputs: torch CUDA tensors or anything np.asarray takes; outputs: float64 / bool CUDA tensors.  There is no CPU path.
"""
import ctypes

import numpy as np

from . import _native as nat
from .host import params_vector


def _dev(device):
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("red_gym_b200.models needs a CUDA device (libf110_b200 has no CPU path)")
    return torch.device("cuda", torch.cuda.current_device() if device is None else int(device))


def _t(v, cols, device):
    import torch
    t = v if isinstance(v, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(v, dtype=np.float64)))
    return t.to(device=device, dtype=torch.float64).reshape(-1, *cols).contiguous()


def _stream(device):
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _p16(params):
    """16 model parameters from a reference params dict (f110_env.py:128) or a sequence in the reference's argument order."""
    v = params_vector(params)[:16] if isinstance(params, dict) else np.asarray(params, dtype=np.float64).reshape(-1)[:16]
    if v.shape[0] != 16:
        raise ValueError("need the 16 vehicle parameters mu .. v_max")
    return np.ascontiguousarray(v, dtype=np.float64)


def _vp(t):
    return ctypes.c_void_p(t.data_ptr())


def dynamics_st(x, u, params, device=None):
    """vehicle_dynamics_st for n states: x (n, 7), u (n, 2) [steering velocity, acceleration] -> f (n, 7)."""
    import torch
    dev = _dev(device)
    x, u = _t(x, (7,), dev), _t(u, (2,), dev)
    f = torch.empty_like(x)
    p = _p16(params)
    nat.check(nat.load().f110_model_dynamics_st(_vp(x), _vp(u), p.ctypes.data_as(ctypes.c_void_p), x.shape[0], _vp(f), dev.index, _stream(dev)))
    return f


def dynamics_ks(x, u, params, device=None):
    """vehicle_dynamics_ks for n states: x (n, 5), u (n, 2) -> f (n, 5)."""
    import torch
    dev = _dev(device)
    x, u = _t(x, (5,), dev), _t(u, (2,), dev)
    f = torch.empty_like(x)
    p = _p16(params)
    nat.check(nat.load().f110_model_dynamics_ks(_vp(x), _vp(u), p.ctypes.data_as(ctypes.c_void_p), x.shape[0], _vp(f), dev.index, _stream(dev)))
    return f


def pid(speed, steer, current_speed, current_steer, max_sv, max_a, max_v, min_v, device=None):
    """pid for n set points -> (accl, sv) tensors of shape (n,)."""
    import torch
    dev = _dev(device)
    cols = [_t(v, (), dev) for v in (speed, steer, current_speed, current_steer)]
    n = max(c.shape[0] for c in cols)
    inp = torch.stack([c.expand(n) for c in cols], dim=1).contiguous()
    out = torch.empty((n, 2), dtype=torch.float64, device=dev)
    nat.check(nat.load().f110_model_pid(_vp(inp), float(max_sv), float(max_a), float(max_v), float(min_v), n, _vp(out), dev.index, _stream(dev)))
    return out[:, 0], out[:, 1]


def _constraint(fn, a, b, limits, device):
    import torch
    dev = _dev(device)
    cols = [_t(v, (), dev) for v in (a, b)]
    n = max(c.shape[0] for c in cols)
    inp = torch.stack([c.expand(n) for c in cols], dim=1).contiguous()
    out = torch.empty((n,), dtype=torch.float64, device=dev)
    nat.check(getattr(nat.load(), fn)(_vp(inp), *[float(v) for v in limits], n, _vp(out), dev.index, _stream(dev)))
    return out


def accl_constraints(vel, accl, v_switch, a_max, v_min, v_max, device=None):
    """accl_constraints (dynamic_models.py:30-60) for n (velocity, acceleration) pairs -> (n,)."""
    return _constraint("f110_model_accl_constraints", vel, accl, (v_switch, a_max, v_min, v_max), device)


def steering_constraint(steering_angle, steering_velocity, s_min, s_max, sv_min, sv_max, device=None):
    """steering_constraint (dynamic_models.py:63-87) for n (angle, velocity) pairs -> (n,)."""
    return _constraint("f110_model_steering_constraint", steering_angle, steering_velocity, (s_min, s_max, sv_min, sv_max), device)


def get_vertices(poses, length, width, device=None):
    """get_vertices for n poses (n, 3) -> corners (n, 4, 2) in the reference's order rl, rr, fr, fl."""
    import torch
    dev = _dev(device)
    poses = _t(poses, (3,), dev)
    v = torch.empty((poses.shape[0], 4, 2), dtype=torch.float64, device=dev)
    nat.check(nat.load().f110_model_vertices(_vp(poses), float(length), float(width), poses.shape[0], _vp(v), dev.index, _stream(dev)))
    return v


def collision(vertices1, vertices2, device=None):
    """GJK overlap of n pairs of convex quads: (n, 4, 2), (n, 4, 2) -> bool (n,)."""
    import torch
    dev = _dev(device)
    a, b = _t(vertices1, (4, 2), dev), _t(vertices2, (4, 2), dev)
    if a.shape[0] != b.shape[0]:
        raise ValueError("vertices1 and vertices2 hold different numbers of bodies")
    hit = torch.empty((a.shape[0],), dtype=torch.uint8, device=dev)
    nat.check(nat.load().f110_model_collision(_vp(a), _vp(b), a.shape[0], _vp(hit), dev.index, _stream(dev)))
    return hit.bool()


def collision_multiple(vertices, device=None):
    """collision_multiple (collision_models.py:185-212) for one set of bodies (A, 4, 2): all pairs i < j on the device
    -> (collisions (A,), collision_idx (A,)) float64 numpy arrays like the reference (last partner wins)."""
    import torch
    v = np.asarray(vertices.cpu() if isinstance(vertices, torch.Tensor) else vertices, dtype=np.float64)
    A = v.shape[0]
    ii, jj = np.triu_indices(A, 1)
    collisions, idx = np.zeros((A,)), -1 * np.ones((A,))
    if len(ii):
        hit = collision(v[ii], v[jj], device).cpu().numpy()
        for i, j in zip(ii[hit], jj[hit]):        # pairs in the reference's visiting order: a later partner overwrites
            collisions[[i, j]] = 1.0
            idx[i], idx[j] = j, i
    return collisions, idx
