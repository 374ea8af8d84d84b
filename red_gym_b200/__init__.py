"""red_gym_b200 -- B200-native batched F1TENTH simulator (hot path of WE-Autopilot/red_gym's f110_gym).

    from red_gym_b200 import BatchF110Env
    env = BatchF110Env(num_envs=4096, map='.../example_map', num_agents=1)

The drop-in single-env API lives in the top-level `f110_gym` package (same import paths as the
reference).  Everything numeric runs in libf110_b200.so (sm_100a CUDA behind a C ABI); importing
this package without that library raises.
"""
from . import _native

_native.load()   # fail loudly when the CUDA library is missing: there is no CPU fallback

from .batch_env import BatchF110Env   # noqa: E402
from .host import DEFAULT_PARAMS, MAPS_DIR, MapData   # noqa: E402
from .lidar_bitmap import LidarBitmap, lidar_to_bitmap   # noqa: E402
from .pure_pursuit import BatchPurePursuitPlanner   # noqa: E402

__all__ = ["BatchF110Env", "DEFAULT_PARAMS", "MAPS_DIR", "MapData", "LidarBitmap", "lidar_to_bitmap", "BatchPurePursuitPlanner"]
