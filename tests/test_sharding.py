"""Multi-GPU host logic on CPU: env partitioning and the statistics all-gather with world_size-2/3 gloo groups.
(`step` itself has no collective; the sharded-vs-unsharded equality of results is a GPU test.)"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from red_gym_b200.sharding import gather_stats, shard_counts, shard_range


@pytest.mark.parametrize("n,world", [(1 << 20, 8), (4096, 2), (10, 3), (7, 8), (1, 1), (65536, 4)])
def test_shard_range_partitions_exactly(n, world):
    ranges = [shard_range(n, r, world) for r in range(world)]
    assert ranges[0][0] == 0
    for (b0, c0), (b1, c1) in zip(ranges, ranges[1:]):
        assert b0 + c0 == b1 and c0 - c1 in (0, 1)
    assert ranges[-1][0] + ranges[-1][1] == n
    assert shard_counts(n, world) == [c for _, c in ranges]
    with pytest.raises(ValueError):
        shard_range(n, world, world)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_total, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        b, c = shard_range(n_total, rank, world)
        # per-env statistics whose value encodes the global env id: the gather must restore global order
        env_ids = torch.arange(b, b + c, dtype=torch.float64)
        local = torch.stack([env_ids, env_ids * 0.5, (env_ids % 2)], dim=-1).reshape(c, 1, 3)
        full = gather_stats(local, n_total)
        ids = torch.arange(n_total, dtype=torch.float64)
        ok = tuple(full.shape) == (n_total, 1, 3) and bool((full[:, 0, 0] == ids).all()) and bool((full[:, 0, 1] == ids * 0.5).all())
        # wrong shard size is rejected
        try:
            gather_stats(local[:-1], n_total) if c > 1 else (_ for _ in ()).throw(ValueError())
            ok = False
        except ValueError:
            pass
        ret[rank] = ok
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_total", [(2, 4096), (3, 10)])
def test_gather_stats_gloo(world, n_total):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(ret.get(r) for r in range(world)), dict(ret)


def test_gather_stats_single_process():
    t = torch.arange(12, dtype=torch.float64).reshape(4, 1, 3)
    assert gather_stats(t, 4) is t
    with pytest.raises(ValueError):
        gather_stats(t, 5)
