"""Env sharding across the GPUs of one box.

Envs are independent (agents interact only inside an env, base_classes.py:570-582), so multi-GPU is pure
partitioning: rank g owns the contiguous env range shard_range(N, g, G), the EDT and tables are replicated
per GPU, and `step` has no collective.  NCCL (NVLink/NVSwitch) is used only when the user asks for the
episode statistics of every env (`gather_stats`), never inside step.
"""
import numpy as np


def shard_range(n_total, rank, world):
    """Contiguous, balanced partition: the first n_total % world ranks get one extra env."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(int(n_total), int(world))
    begin = rank * base + min(rank, rem)
    count = base + (1 if rank < rem else 0)
    return begin, count


def shard_counts(n_total, world):
    return [shard_range(n_total, r, world)[1] for r in range(world)]


STAT_FIELDS = ("lap_count", "lap_time", "collided", "sim_time", "episodes")


def episode_stats(env):
    """(count, A, 5) float64 tensor on the env's device: lap_counts, lap_times, collisions, current time, episodes."""
    import torch
    N, A = env.num_envs, env.num_agents
    t = env.current_time.reshape(N, 1).expand(N, A)
    ep = env.episode_count.to(torch.float64).reshape(N, 1).expand(N, A)
    return torch.stack([env.lap_counts, env.lap_times, env.collisions, t, ep], dim=-1).contiguous()


def gather_stats(local, n_total, group=None):
    """all_gather of per-env statistics over the ranks of the default process group.

    local: (count_r, ...) tensor of this rank's shard (CUDA -> NCCL over NVLink, CPU -> gloo).  Returns the
    (n_total, ...) tensor in global env order on every rank.  Shards may differ by one env: they are padded to
    the largest shard for the collective and trimmed afterwards."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        if local.shape[0] != n_total:
            raise ValueError("single process: local shard must hold every env")
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    counts = shard_counts(n_total, world)
    if local.shape[0] != counts[rank]:
        raise ValueError("rank %d holds %d envs, expected %d" % (rank, local.shape[0], counts[rank]))
    cmax = max(counts)
    pad = torch.zeros((cmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:counts[rank]] = local
    out = torch.empty((world * cmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    out = out.reshape((world, cmax) + tuple(local.shape[1:]))
    return torch.cat([out[r, :counts[r]] for r in range(world)], dim=0)


class ShardedF110Env(object):
    """This rank's shard of an N-env batch.  Same API as BatchF110Env for the local envs; `global_slice` tells
    which envs they are; `gather_stats()` returns the statistics of all N envs."""

    def __init__(self, num_envs_total, rank=None, world=None, **kwargs):
        import os
        from .batch_env import BatchF110Env
        rank = int(os.environ.get("RANK", 0)) if rank is None else rank
        world = int(os.environ.get("WORLD_SIZE", 1)) if world is None else world
        local = int(os.environ.get("LOCAL_RANK", rank))
        self.n_total, self.rank, self.world = int(num_envs_total), rank, world
        self.begin, self.count = shard_range(num_envs_total, rank, world)
        kwargs.setdefault("device", local)
        kwargs["env_offset"] = self.begin
        self.env = BatchF110Env(num_envs=self.count, **kwargs)

    @property
    def global_slice(self):
        return slice(self.begin, self.begin + self.count)

    def reset(self, poses_global, mask_global=None):
        poses = np.asarray(poses_global)[self.global_slice]
        mask = None if mask_global is None else np.asarray(mask_global)[self.global_slice]
        return self.env.reset(poses, mask=mask)

    def step(self, actions_local):
        return self.env.step(actions_local)

    def gather_stats(self):
        return gather_stats(episode_stats(self.env), self.n_total)

    def __getattr__(self, name):
        return getattr(self.env, name)
