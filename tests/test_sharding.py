"""The N > 1 path on CPU: two gloo ranks shard a job, step their envs with the oracle (standing in
for the GPU only as far as producing per-shard accumulators), and all-reduce the episode stats."""

import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from metta_b200.sharding import reduce_episode_accumulators, shard_maps, shard_range, shard_seeds


def test_shards_tile_the_job_and_do_not_depend_on_world_size():
    total, pool = 64, 5
    ref_seeds = shard_seeds(0, 1, total, 42)
    ref_maps = shard_maps(0, 1, total, pool)
    for world in (2, 4, 8):
        per = total // world
        seeds = np.concatenate([shard_seeds(r, world, per, 42) for r in range(world)])
        maps = np.concatenate([shard_maps(r, world, per, pool) for r in range(world)])
        assert np.array_equal(seeds, ref_seeds) and np.array_equal(maps, ref_maps)
        covered = [i for r in range(world) for i in shard_range(r, world, per)]
        assert covered == list(range(total))


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    per = 3
    seeds = shard_seeds(rank, world, per, 100)
    # a per-shard accumulator like Batch.read_episode_accum(): episodes, reward sum, then stat slots
    acc = torch.zeros(10, dtype=torch.float64)
    acc[0] = per
    acc[1] = float(seeds.sum())
    acc[2 + rank] = 1.0
    reduce_episode_accumulators(acc)
    if rank == 0:
        out.put(acc.numpy().copy())
    dist.barrier()
    dist.destroy_process_group()


def test_episode_stats_all_reduce_over_two_gloo_ranks():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    acc = out.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert acc[0] == 6
    assert acc[1] == float(sum(range(100, 106)))
    assert acc[2] == 1.0 and acc[3] == 1.0
