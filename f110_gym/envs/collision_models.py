"""Drop-in for the reference's f110_gym.envs.collision_models (gym/f110_gym/envs/collision_models.py:114-260):
`collision` (GJK), `collision_multiple` and `get_vertices` with the reference's signatures, evaluated by the B200
library (red_gym_b200.models).  One call = one small kernel launch; use red_gym_b200.models for batches."""
import numpy as np

from red_gym_b200 import models as _m


def collision(vertices1, vertices2):
    """GJK overlap test of two convex quads (4, 2) -> bool."""
    return bool(_m.collision(np.asarray(vertices1, dtype=np.float64)[None], np.asarray(vertices2, dtype=np.float64)[None])[0])


def collision_multiple(vertices):
    """Pair-wise GJK over (num_bodies, 4, 2) -> (collisions, collision_idx), both (num_bodies,) float64."""
    return _m.collision_multiple(vertices)


def get_vertices(pose, length, width):
    """Corners of the car rectangle at pose (x, y, theta) -> (4, 2): rear-left, rear-right, front-right, front-left."""
    return _m.get_vertices(np.asarray(pose, dtype=np.float64)[None, :3], length, width)[0].cpu().numpy()
