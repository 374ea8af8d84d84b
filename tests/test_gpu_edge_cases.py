"""Edge sizes of the domain on the GPU, against the CPU oracle: many agents per env (up to F110_MAX_AGENTS = 32: 496 GJK pairs
and 31 opponent casts per agent, collision_idx last-writer-wins), and beam counts that hit every shape of the kernels' beam
loops -- fewer beams than a warp, exactly one 256-beam block, a block plus a tail, the 65 535-beam maximum."""
import os

import numpy as np
import pytest

import f110_oracle as fo
from conftest import MAPS

pytestmark = pytest.mark.gpu
REL, SCAN_TOL = 1e-9, 1e-4


def _close(a, b, rel=REL):
    return np.abs(a - b) <= rel * np.maximum(1.0, np.abs(b))


@pytest.mark.parametrize("A,N,T", [(8, 64, 40), (32, 6, 10)])
def test_many_agents_rollout_vs_oracle(A, N, T):
    from red_gym_b200 import BatchF110Env
    from red_gym_b200.host import raceline_poses
    rng = np.random.default_rng(A)
    poses = np.repeat(raceline_poses(N)[:, None, :], A, axis=1)
    poses[:, 1:, 0] += rng.uniform(-2.5, 2.5, (N, A - 1))
    poses[:, 1:, 1] += rng.uniform(-2.5, 2.5, (N, A - 1))
    poses[:, 1:, 2] = rng.uniform(-3, 3, (N, A - 1))
    kw = dict(map=os.path.join(MAPS, "example_map"), seed=7, ego_idx=A - 1)
    env = BatchF110Env(num_envs=N, num_agents=A, **kw)
    ora = fo.OracleEnv(N, A, noise_rows=T + 8, **kw)
    env.reset(poses); ora.reset(poses)
    bad = np.zeros(N, dtype=bool)
    hits = 0
    for t in range(T + 1):
        if t > 0:
            act = np.stack([rng.uniform(-0.3, 0.3, (N, A)), rng.uniform(0.0, 5.0, (N, A))], axis=-1)
            env.step(act); ora.step(act)
        st = env.state.permute(1, 2, 0).cpu().numpy()
        ok = (env.collisions.cpu().numpy() == ora.collisions).all(axis=1) & (env.collision_idx.cpu().numpy() == ora.collision_idx).all(axis=1) & \
             (env.toggle_list.cpu().numpy() == ora.toggles).all(axis=1) & (env.done.cpu().numpy().astype(bool) == ora.done.astype(bool)) & \
             _close(st, ora.state).all(axis=(1, 2))
        # opponent casts: a beam grazing a corner may fall on the other side in the last ulp of sin / cos
        off = np.abs(env.scans.cpu().numpy() - ora.scans) > SCAN_TOL
        ok &= off.mean(axis=(1, 2)) < 1e-3
        bad |= ~ok
        hits += int((ora.collision_idx >= 0).sum())
    assert bad.sum() <= (1 if N >= 32 else 0), "envs diverged from the oracle: %s" % np.flatnonzero(bad)
    assert hits > 0                                     # GJK did fire


@pytest.mark.parametrize("B,fov,n", [(2, 4.7, 300), (3, 2 * np.pi, 300), (31, 4.7, 300), (33, 2 * np.pi, 300), (256, 4.7, 300), (65535, 2 * np.pi, 24)])
def test_raycast_beam_count_edges(B, fov, n):
    from red_gym_b200 import BatchF110Env
    ora = fo.OracleEnv(1, 1, map=os.path.join(MAPS, "berlin"), noise=False, num_beams=B, fov=fov)
    rng = np.random.default_rng(B)
    poses = np.stack([rng.uniform(-0.4, 0.4, n), rng.uniform(-0.4, 0.4, n), rng.uniform(-7, 7, n)], axis=1)
    poses[0] = [1e3, -1e3, 0.5]                         # outside the map
    want = ora.scan(poses)
    kw = dict(num_envs=1, num_agents=1, map=os.path.join(MAPS, "berlin"), num_beams=B, fov=fov, noise="off")
    for variant in (0, 128):                            # CTA-per-scan and warp-per-scan kernels
        got = BatchF110Env(dtype="fp64", raycast_variant=variant, **kw).raycast(poses).cpu().numpy()
        assert (got == want).all(), (B, variant, np.abs(got - want).max())
    g32 = [BatchF110Env(dtype="fp32", raycast_variant=v, **kw).raycast(poses).cpu().numpy() for v in (0, 128)]
    assert (g32[0] == g32[1]).all()
    d = np.abs(g32[0] - want)
    assert np.median(d) < 1e-3 and (d > 2e-2).mean() < 0.03


@pytest.mark.parametrize("B", [33, 256, 300])
def test_step_with_other_beam_counts(B):
    """The whole step at beam counts other than 1080: fp64 against the oracle (iTTC tables, noise rows and the fused /
    split / whole-scan kernels all depend on B), fp32 CTA-per-scan against warp-per-scan, and the uint16 format."""
    from red_gym_b200 import BatchF110Env
    from red_gym_b200.host import raceline_poses
    import torch
    N, T = 400, 160                                     # long enough for cars that start at rest to reach a wall
    poses = raceline_poses(N)[:, None, :]
    kw = dict(map=os.path.join(MAPS, "example_map"), num_beams=B, seed=3)
    rng = np.random.default_rng(B)
    acts = np.stack([rng.uniform(-0.4, 0.4, (T, N, 1)), rng.uniform(0.0, 8.0, (T, N, 1))], axis=-1)
    ora = fo.OracleEnv(N, 1, noise_rows=T + 8, **kw)
    envs = [BatchF110Env(num_envs=N, num_agents=1, raycast_variant=v, **kw) for v in (0, 128)]
    f32 = [BatchF110Env(num_envs=N, num_agents=1, dtype="fp32", noise="philox", raycast_variant=v, scan_format=f, **kw)
           for v, f in ((0, "native"), (128, "native"), (0, "u16" if B % 2 == 0 else "native"))]
    for e in envs + f32 + [ora]:
        e.reset(poses)
    bad = np.zeros(N, dtype=bool)
    ever = False
    for t in range(T):
        for e in envs + f32 + [ora]:
            e.step(acts[t])
        ever |= bool(ora.collisions.any())
        st = envs[0].state.permute(1, 2, 0).cpu().numpy()
        bad |= ~((envs[0].collisions.cpu().numpy() == ora.collisions).all(axis=1) & _close(st, ora.state).all(axis=(1, 2)) &
                 (np.abs(envs[0].scans.cpu().numpy() - ora.scans).max(axis=(1, 2)) <= SCAN_TOL))
    assert bad.sum() <= 1, np.flatnonzero(bad)
    assert ever or B < 64          # iTTC fired somewhere in the rollout (33 beams, 11 degrees apart, may miss every wall contact)
    assert torch.equal(envs[0].obs_block, envs[1].obs_block) and torch.equal(envs[0].scans, envs[1].scans)
    assert torch.equal(f32[0].obs_block, f32[1].obs_block) and torch.equal(f32[0].scans, f32[1].scans)
    assert torch.equal(f32[0].obs_block, f32[2].obs_block)
    if B % 2 == 0:
        q = f32[2].scans.to(torch.int32).double() / 2000.0
        assert float((q - f32[0].scans.double().clamp(min=0)).abs().max()) <= 0.25e-3 + 2e-6


@pytest.mark.parametrize("dtype,N", [("fp32", 16384), ("fp64", 20480)])
def test_sort_thresholds_per_precision(dtype, N):
    """The spatial sort starts at F110_ORDER_FROM32 = 16384 (fp32) / F110_ORDER_FROM64 = 20480 (fp64) scans (measured break-even,
    f110_dev.h): on at exactly that size (three more launches per step than raycast_variant=64), off one scan below, and only a
    schedule -- step and step_host (8 copy-out chunks of 2048 / 2560 envs, chunk-major order) stay bit-identical."""
    from red_gym_b200 import BatchF110Env
    from red_gym_b200.host import raceline_poses
    import torch
    kw = dict(num_agents=1, map=os.path.join(MAPS, "example_map"), dtype=dtype, noise="philox" if dtype == "fp32" else "numpy",
              auto_reset=True, reset_poses=raceline_poses(128)[:, None, :])
    a, b = BatchF110Env(num_envs=N, **kw), BatchF110Env(num_envs=N, raycast_variant=64, **kw)
    below = BatchF110Env(num_envs=N - 1, **kw)
    rng = np.random.default_rng(1)
    poses = raceline_poses(N)[rng.permutation(N)][:, None, :]
    a.reset(poses); b.reset(poses); below.reset(poses[:N - 1])
    l0 = (a.launch_count, b.launch_count, below.launch_count)
    g = torch.Generator(device="cuda").manual_seed(4)
    sdt = torch.float32 if dtype == "fp32" else torch.float64
    hb = [dict(obs=torch.zeros((10, N, 1), dtype=torch.float64).pin_memory(), scans=torch.zeros((N, 1, 1080), dtype=sdt).pin_memory(),
               done=torch.zeros((N,), dtype=torch.uint8).pin_memory()) for _ in range(2)]
    T = 30
    for t in range(T):
        act = torch.rand((N, 1, 2), generator=g, device="cuda", dtype=torch.float64)
        act[..., 0] = (act[..., 0] * 2 - 1) * 0.4189
        act[..., 1] = act[..., 1] * 8.0
        if t % 10 == 9:
            ah = act.cpu().pin_memory()
            a.step_host(ah, hb[0]["obs"], hb[0]["scans"], hb[0]["done"]); b.step_host(ah, hb[1]["obs"], hb[1]["scans"], hb[1]["done"])
            assert all(torch.equal(hb[0][k], hb[1][k]) for k in ("obs", "scans", "done")) and torch.equal(hb[0]["scans"].cuda(), a.scans)
        else:
            a.step(act); b.step(act)
        below.step(act[:N - 1].contiguous())
        assert torch.equal(a.obs_block, b.obs_block) and torch.equal(a.scans, b.scans), t
    steps_dev = T - T // 10
    assert (a.launch_count - l0[0]) - (b.launch_count - l0[1]) == 3 * T          # the three sort kernels, every step
    assert below.launch_count - l0[2] == 3 * T                                   # one scan below the threshold: no sort
    assert torch.equal(below.obs_block, a.obs_block[:, :N - 1])                  # ... and the same results


def test_opponent_cast_with_a_yaw_far_outside_one_turn():
    """The reference wraps the yaw by 2 pi once per step only (base_classes.py:398-402) and casts opponents along
    pose[2] + scan_angles[i] as rounded (laser_models.py:343): at a yaw of 1e15 doubles are 0.125 rad apart and neighbouring
    beams share a direction.  The library's beam filter (cone around the opponent, edge-parallel beams) assumes exact
    directions and must step aside there (found by tools/step_fuzz.py: a car whose model blew up under a harsh parameter set)."""
    from red_gym_b200 import BatchF110Env
    from red_gym_b200.host import raceline_poses
    N, A = 240, 3
    rng = np.random.default_rng(5)
    poses = np.repeat(raceline_poses(N)[:, None, :], A, axis=1)
    poses[:, 1:, 0] += rng.uniform(-2.0, 2.0, (N, A - 1))
    poses[:, 1:, 1] += rng.uniform(-2.0, 2.0, (N, A - 1))
    big = np.array([5.9028286868952100e+14, -3.3e12, 1.0e7, 7.7e9, 2.5e15, -8.1e13])
    poses[:, :, 2] = rng.uniform(-3, 3, (N, A))
    poses[:, 0, 2] = big[np.arange(N) % big.size] + rng.uniform(-3, 3, N)          # the scanning car's yaw
    poses[N // 2:, 1, 2] = big[np.arange(N - N // 2) % big.size]                     # an opponent's yaw
    kw = dict(map=os.path.join(MAPS, "example_map"), seed=9, num_beams=540, fov=4.7)
    for dtype in ("fp64", "fp32"):
        env = BatchF110Env(num_envs=N, num_agents=A, dtype=dtype, noise="off" if dtype == "fp32" else "numpy", **kw)
        ora = fo.OracleEnv(N, A, noise_rows=8, noise=(dtype == "fp64"), **kw)
        env.reset(poses); ora.reset(poses)
        changed = 0
        for t in range(3):
            act = np.stack([rng.uniform(-0.3, 0.3, (N, A)), rng.uniform(0.0, 3.0, (N, A))], axis=-1)
            env.step(act); ora.step(act)
            off = np.abs(env.scans.cpu().numpy() - ora.scans) > (SCAN_TOL if dtype == "fp64" else 2e-2)
            # fp64: a beam may graze a corner differently in the last ulp of sin / cos of a huge argument; whole groups of beams may not
            assert off.sum(axis=2).max() <= (2 if dtype == "fp64" else 12), (dtype, t, np.argwhere(off.sum(axis=2) > 2)[:5])
            assert (env.collision_idx.cpu().numpy() == ora.collision_idx).all()
            changed += int((ora.scans < 29.0).sum())
        assert changed > 0
