"""Drop-in for the reference's f110_gym.envs.dynamic_models (gym/f110_gym/envs/dynamic_models.py:30-221): the same
function names, argument lists and return shapes, evaluated by the B200 library (red_gym_b200.models -> the device
functions the step kernel uses).  One call = one small kernel launch; use red_gym_b200.models for batches."""
import numpy as np

from red_gym_b200 import models as _m


def _one(t):
    return t.cpu().numpy()[0]


def vehicle_dynamics_ks(x, u_init, mu, C_Sf, C_Sr, lf, lr, h, m, I, s_min, s_max, sv_min, sv_max, v_switch, a_max, v_min, v_max):
    """Kinematic single-track right-hand side: x (5,), u_init (2,) -> f (5,)."""
    p = [mu, C_Sf, C_Sr, lf, lr, h, m, I, s_min, s_max, sv_min, sv_max, v_switch, a_max, v_min, v_max]
    return _one(_m.dynamics_ks(np.asarray(x, dtype=np.float64)[None, :5], np.asarray(u_init, dtype=np.float64)[None, :2], p))


def vehicle_dynamics_st(x, u_init, mu, C_Sf, C_Sr, lf, lr, h, m, I, s_min, s_max, sv_min, sv_max, v_switch, a_max, v_min, v_max):
    """Single-track right-hand side (kinematic below 0.5 m/s): x (7,), u_init (2,) -> f (7,)."""
    p = [mu, C_Sf, C_Sr, lf, lr, h, m, I, s_min, s_max, sv_min, sv_max, v_switch, a_max, v_min, v_max]
    return _one(_m.dynamics_st(np.asarray(x, dtype=np.float64)[None, :7], np.asarray(u_init, dtype=np.float64)[None, :2], p))


def accl_constraints(vel, accl, v_switch, a_max, v_min, v_max):
    """Acceleration limits (dynamic_models.py:30-60): the step kernel's own device function (f110_model_accl_constraints)."""
    return float(_one(_m.accl_constraints(vel, accl, v_switch, a_max, v_min, v_max)))


def steering_constraint(steering_angle, steering_velocity, s_min, s_max, sv_min, sv_max):
    """Steering limits (dynamic_models.py:63-87): the step kernel's own device function (f110_model_steering_constraint)."""
    return float(_one(_m.steering_constraint(steering_angle, steering_velocity, s_min, s_max, sv_min, sv_max)))


def pid(speed, steer, current_speed, current_steer, max_sv, max_a, max_v, min_v):
    """Speed / steer set points -> (accl, sv)."""
    a, sv = _m.pid(speed, steer, current_speed, current_steer, max_sv, max_a, max_v, min_v)
    return float(a[0]), float(sv[0])


def func_KS(x, t, u, mu, C_Sf, C_Sr, lf, lr, h, m, I, s_min, s_max, sv_min, sv_max, v_switch, a_max, v_min, v_max):
    """scipy.integrate.odeint signature of the kinematic model."""
    return vehicle_dynamics_ks(x, u, mu, C_Sf, C_Sr, lf, lr, h, m, I, s_min, s_max, sv_min, sv_max, v_switch, a_max, v_min, v_max)


def func_ST(x, t, u, mu, C_Sf, C_Sr, lf, lr, h, m, I, s_min, s_max, sv_min, sv_max, v_switch, a_max, v_min, v_max):
    """scipy.integrate.odeint signature of the single-track model."""
    return vehicle_dynamics_st(x, u, mu, C_Sf, C_Sr, lf, lr, h, m, I, s_min, s_max, sv_min, sv_max, v_switch, a_max, v_min, v_max)
