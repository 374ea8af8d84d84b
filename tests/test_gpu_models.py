"""The reference's own unit tests of the model functions (DynamicsTest in dynamic_models.py:231-423, CollisionTests in
collision_models.py:262-324), run against the device code through the C ABI (f110_model_*), plus the vectors the
unmodified reference produced for tests/golden/kats.npz.  These are the device functions the step kernels call."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
G = 9.81
# DynamicsTest.setUp (dynamic_models.py:232-253): the full-size vehicle of the CommonRoad models
VEH = dict(mu=1.0489, C_Sf=21.92 / 1.0489, C_Sr=21.92 / 1.0489, lf=0.3048 * 3.793293, lr=0.3048 * 4.667707, h=0.3048 * 2.01355,
           m=4.4482216152605 / 0.3048 * 74.91452, I=4.4482216152605 * 0.3048 * 1321.416, s_min=-1.066, s_max=1.066,
           sv_min=-0.4, sv_max=0.4, v_switch=7.319, a_max=11.5, v_min=-13.6, v_max=50.8)
ORDER = ["mu", "C_Sf", "C_Sr", "lf", "lr", "h", "m", "I", "s_min", "s_max", "sv_min", "sv_max", "v_switch", "a_max", "v_min", "v_max"]
ARGS = [VEH[k] for k in ORDER]


def test_derivatives():
    """test_derivatives (:255-278): published right-hand sides of both models, through the drop-in module."""
    from f110_gym.envs.dynamic_models import vehicle_dynamics_ks, vehicle_dynamics_st
    f_ks_gt = [16.3475935934250209, 0.4819314886013121, 0.1500000000000000, 5.1464424102339752, 0.2401426578627629]
    f_st_gt = [15.7213512030862397, 0.0925527979719355, 0.1500000000000000, 5.3536773276413925, 0.0529001056654038,
               0.6435589397748606, 0.0313297971641291]
    x_ks = np.array([3.9579422297936526, 0.0391650102771405, 0.0378491427211811, 16.3546957860883566, 0.0294717351052816])
    x_st = np.array([2.0233348142065677, 0.0041907137716636, 0.0197545248559617, 15.7216236334290116, 0.0025857914776859,
                     0.0529001056654038, 0.0033012170610298])
    u = np.array([0.15, 0.63 * G])
    f_ks = vehicle_dynamics_ks(x_ks, u, *ARGS)
    f_st = vehicle_dynamics_st(x_st, u, *ARGS)
    assert f_ks.shape == (5,) and f_st.shape == (7,)
    assert np.max(np.abs(np.array(f_ks_gt) - f_ks)) < 5e-8          # assertAlmostEqual: 7 places
    assert np.max(np.abs(np.array(f_st_gt) - f_st)) < 5e-8


def _integrate(u, steps=1000, h=1e-3):
    """RK4 from rest for 1 s on the device (batch of one state, both models side by side)."""
    import torch
    from red_gym_b200 import models
    x = torch.zeros((1, 7), dtype=torch.float64, device="cuda")
    k = torch.zeros((1, 5), dtype=torch.float64, device="cuda")
    uu = torch.tensor([u], dtype=torch.float64, device="cuda")
    for _ in range(steps):
        k1 = models.dynamics_st(x, uu, ARGS); k2 = models.dynamics_st(x + h * k1 / 2, uu, ARGS)
        k3 = models.dynamics_st(x + h * k2 / 2, uu, ARGS); k4 = models.dynamics_st(x + h * k3, uu, ARGS)
        x = x + h / 6 * (k1 + 2 * k2 + 2 * k3 + k4)
        q1 = models.dynamics_ks(k, uu, ARGS); q2 = models.dynamics_ks(k + h * q1 / 2, uu, ARGS)
        q3 = models.dynamics_ks(k + h * q2 / 2, uu, ARGS); q4 = models.dynamics_ks(k + h * q3, uu, ARGS)
        k = k + h / 6 * (q1 + 2 * q2 + 2 * q3 + q4)
    return x[0].cpu().numpy(), k[0].cpu().numpy()


def test_zeroinit_roll_dec_acc_rollleft():
    """test_zeroinit_* (:281-423): 1 s from rest under four inputs, published end states within 1e-2."""
    st, ks = _integrate([0.0, 0.0])
    assert (st == 0).all() and (ks == 0).all()                                                       # :310-311
    st, ks = _integrate([0.0, -0.7 * G])
    assert np.all(np.abs(st - [-3.4335000000000013, 0., 0., -6.8670000000000018, 0., 0., 0.]) < 1e-2)       # :343
    assert np.all(np.abs(ks - [-3.4335000000000013, 0., 0., -6.8670000000000018, 0.]) < 1e-2)                # :344
    st, ks = _integrate([0.15, 0.63 * G])
    assert np.all(np.abs(st - [3.0731976046859715, 0.2869835398304389, 0.15, 6.1802999999999999, 0.1097747074946325,
                               0.3248268063223301, 0.0697547542798040]) < 1e-2)                               # :381
    assert np.all(np.abs(ks - [3.0845676868494927, 0.1484249221523042, 0.15, 6.1803000000000017, 0.1203664469224163]) < 1e-2)
    st, ks = _integrate([0.15, 0.0])
    assert np.all(np.abs(st - [0., 0., 0.15, 0., 0., 0., 0.]) < 1e-2)                                         # :418
    assert np.all(np.abs(ks - [0., 0., 0.15, 0., 0.]) < 1e-2)


def test_reference_vectors_dynamics_pid_constraints():
    """Random states / set points evaluated by the unmodified reference (tests/golden/kats.npz)."""
    from f110_gym.envs.dynamic_models import accl_constraints, pid, steering_constraint
    from red_gym_b200 import models
    k = np.load(os.path.join(GOLDEN, "kats.npz"))
    f = models.dynamics_st(k["dyn_x"], k["dyn_u"], k["dyn_params"]).cpu().numpy()
    assert np.abs(f - k["dyn_f"]).max() < 1e-12                       # CUDA sin/cos/tan vs glibc: last bits
    a, sv = models.pid(k["pid_in"][:, 0], k["pid_in"][:, 1], k["pid_in"][:, 2], k["pid_in"][:, 3], 3.2, 9.51, 20.0, -5.0)
    assert (a.cpu().numpy() == k["pid_out"][:, 0]).all() and (sv.cpu().numpy() == k["pid_out"][:, 1]).all()
    assert pid(*k["pid_in"][0], 3.2, 9.51, 20.0, -5.0) == (k["pid_out"][0, 0], k["pid_out"][0, 1])
    # the two constraint helpers (dynamic_models.py:29-87), corner cases spelled out
    assert accl_constraints(8.0, 20.0, 7.319, 9.51, -5.0, 20.0) == 9.51 * 7.319 / 8.0
    assert accl_constraints(1.0, 20.0, 7.319, 9.51, -5.0, 20.0) == 9.51 and accl_constraints(1.0, -20.0, 7.319, 9.51, -5.0, 20.0) == -9.51
    assert accl_constraints(20.0, 1.0, 7.319, 9.51, -5.0, 20.0) == 0.0 and accl_constraints(-5.0, -1.0, 7.319, 9.51, -5.0, 20.0) == 0.0
    assert steering_constraint(0.4189, 1.0, -0.4189, 0.4189, -3.2, 3.2) == 0.0 and steering_constraint(0.0, 5.0, -0.4189, 0.4189, -3.2, 3.2) == 3.2
    assert steering_constraint(0.0, -5.0, -0.4189, 0.4189, -3.2, 3.2) == -3.2 and steering_constraint(0.1, 0.7, -0.4189, 0.4189, -3.2, 3.2) == 0.7


def test_get_vertices_and_collision():
    """CollisionTests (:262-324): get_vertices against the reference's output, the 1000 jittered overlapping quads
    of test_random_collision, and test_multiple_collisions' known answer under np.random.seed(1234)."""
    from f110_gym.envs.collision_models import collision, collision_multiple, get_vertices
    from red_gym_b200 import models
    k = np.load(os.path.join(GOLDEN, "kats.npz"))
    va = models.get_vertices(k["car_poses_a"], 0.58, 0.31)
    vb = models.get_vertices(k["car_poses_b"], 0.58, 0.31)
    assert np.abs(va.cpu().numpy() - k["car_verts_a"]).max() < 1e-14 and np.abs(vb.cpu().numpy() - k["car_verts_b"]).max() < 1e-14
    assert (models.collision(va, vb).cpu().numpy() == k["car_hit"]).all()
    assert np.abs(get_vertices(k["car_poses_a"][0], 0.58, 0.31) - k["car_verts_a"][0]).max() < 1e-14
    assert (models.collision(k["gjk_pairs_a"], k["gjk_pairs_b"]).cpu().numpy() == k["gjk_pairs_hit"]).all()
    np.random.seed(1234)
    v1 = np.asarray([[4, 11.], [5, 5], [9, 9], [10, 10]])
    a = np.stack([v1 + np.random.normal(size=v1.shape) / 100. for _ in range(2000)])
    assert models.collision(a[0::2], a[1::2]).all()                  # test_random_collision (:306-311)
    assert collision(a[0], a[1]) is True
    c, ci = collision_multiple(k["gjk_bodies"])
    assert (c == np.array([1., 1., 1., 1., 1., 1., 0.])).all()       # :320-324
    assert (ci == np.array([5., 5., 5., 5., 5., 4., -1.])).all()
