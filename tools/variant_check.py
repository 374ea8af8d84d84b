"""A/B helper: scans of raycast_variant V must equal variant 0 bit for bit (python tools/variant_check.py V)."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from red_gym_b200 import BatchF110Env, MAPS_DIR
from red_gym_b200.host import raceline_poses
V = int(sys.argv[1]); N = 20000
poses = raceline_poses(N)[:, None, :]
kw = dict(num_envs=N, num_agents=1, dtype="fp32", noise="philox", map=os.path.join(MAPS_DIR, "example_map"))
a = BatchF110Env(**kw); b = BatchF110Env(raycast_variant=V, **kw)
a.reset(poses); b.reset(poses)
g = torch.Generator(device="cuda").manual_seed(1)
for t in range(30):
    act = torch.rand((N, 1, 2), generator=g, device="cuda", dtype=torch.float64); act[..., 0] = (act[..., 0] * 2 - 1) * 0.4189; act[..., 1] *= 8
    a.step(act); b.step(act)
print("variant", V, "identical:", bool((a.scans == b.scans).all()), bool((a.obs_block == b.obs_block).all()))
