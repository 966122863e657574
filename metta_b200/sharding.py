"""Env shards across GPUs: one process per GPU, envs independent, no data-path collective.

SURVEY.md 8(e): rank r of W owns the global env indices [r*N, (r+1)*N). Seeds and map indices
derive from the GLOBAL index, so what an env does never depends on how many GPUs share the job.
The only collective is a SUM all-reduce of the finished-episode accumulators (a few KB).
"""

from __future__ import annotations

import numpy as np


def shard_range(rank: int, world: int, envs_per_rank: int) -> range:
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    return range(rank * envs_per_rank, (rank + 1) * envs_per_rank)


def shard_seeds(rank: int, world: int, envs_per_rank: int, base_seed: int) -> np.ndarray:
    idx = np.asarray(shard_range(rank, world, envs_per_rank), dtype=np.int64)
    return ((idx + int(base_seed)) & 0xFFFFFFFF).astype(np.uint32)


def shard_maps(rank: int, world: int, envs_per_rank: int, pool_size: int) -> np.ndarray:
    idx = np.asarray(shard_range(rank, world, envs_per_rank), dtype=np.int64)
    return (idx % int(pool_size)).astype(np.int32)


def reduce_episode_accumulators(acc, group=None):
    """SUM all-reduce of the accumulator vector [episodes, reward sum, agent stats..., game stats...].
    `acc` is a float64 torch tensor (CUDA with NCCL, CPU with gloo). Returns it, reduced in place."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(acc, op=dist.ReduceOp.SUM, group=group)
    return acc
