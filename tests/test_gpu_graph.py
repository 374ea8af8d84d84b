"""f110_step under CUDA graph capture: the step enqueues kernel launches only (no allocation, no synchronisation), so it can be
captured once and replayed with one launch per replay.  Replays must give the bits of eager steps -- on the fused one-kernel
path (a handful of fp64 envs), the CTA-per-scan path, and the spatially sorted large-launch path (three sort kernels, scan
records, cudaFuncSetAttribute inside the capture)."""
import os

import numpy as np
import pytest

from conftest import MAPS

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype,N,A", [("fp64", 8, 1), ("fp64", 700, 1), ("fp32", 3000, 1), ("fp32", 40000, 1), ("fp64", 512, 2)])
def test_graph_replay_is_bit_identical_to_eager_steps(dtype, N, A):
    import torch
    from red_gym_b200 import BatchF110Env
    from red_gym_b200.host import raceline_poses
    lead = raceline_poses(N)
    poses = [lead]
    for a in range(1, A):
        c = lead.copy()
        c[:, 0] -= 1.5 * a * np.cos(lead[:, 2]); c[:, 1] -= 1.5 * a * np.sin(lead[:, 2])
        poses.append(c)
    poses = np.stack(poses, axis=1)
    pool = poses[:64].copy()
    kw = dict(num_envs=N, num_agents=A, map=os.path.join(MAPS, "example_map"), dtype=dtype, auto_reset=True, reset_poses=pool,
              noise="philox" if dtype == "fp32" else "numpy")
    eager, graphed = BatchF110Env(**kw), BatchF110Env(**kw)
    eager.reset(poses); graphed.reset(poses)
    g = torch.Generator(device="cuda").manual_seed(6)
    K, R = 4, 30                                 # K steps per graph, R replays
    bank = torch.rand((R, K, N, A, 2), generator=g, device="cuda", dtype=torch.float64)
    bank[..., 0] = (bank[..., 0] * 2 - 1) * 0.4189
    bank[..., 1] = bank[..., 1] * 8.0
    static = bank[0].clone()                     # the graph reads this tensor; it is overwritten in place before every replay
    warm = 3
    graphed.reserve_noise(warm + K * R + 2)
    graph = graphed.capture_step_graph(static, steps=K, warmup=warm)     # takes `warm` eager steps with static[k % K]
    for k in range(warm):
        eager.step(static[k % K])
    l0 = graphed.launch_count
    for r in range(R):
        static.copy_(bank[r])
        graph.replay()
        for k in range(K):
            eager.step(bank[r, k])
        if r % 10 == 9 or r == R - 1:
            assert torch.equal(graphed.obs_block, eager.obs_block) and torch.equal(graphed.scans, eager.scans), r
            assert torch.equal(graphed.done, eager.done) and torch.equal(graphed.episode_count, eager.episode_count)
            assert torch.equal(graphed.noise_step, eager.noise_step) and torch.equal(graphed.collision_idx, eager.collision_idx)
    assert graphed.launch_count == l0            # replays go through the graph, not through the library's launch path
    assert int(eager.episode_count.max()) >= 2 or N < 64
    graphed.note_replayed(K * R)
    a = bank[0, 0]
    graphed.step(a); eager.step(a)               # and eager stepping carries on from there
    assert torch.equal(graphed.obs_block, eager.obs_block) and torch.equal(graphed.scans, eager.scans)
