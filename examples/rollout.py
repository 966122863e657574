"""BASELINE.json configs[4] in small: a rollout loop with a torch policy in it.

    python examples/rollout.py [--envs 4096] [--steps 200]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/rollout.py

The env writes observations straight into the tensor the policy reads (no host copy, no extra
device copy), the policy's sampled actions go back as a device tensor, and everything is ordered on
one CUDA stream. The policy is a stand-in of the reference's token pipeline (strip 0xFF padding,
embed (feature, value) per token, pool, act: agent/src/metta/agent/components/obs_shim.py:166-190),
not a trained model. One process per GPU; envs are sharded by rank (metta_b200/sharding.py) and the
only collective is the all-reduce of episode statistics at the end.
"""

import argparse
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from metta_b200.builders import make_arena  # noqa: E402
from metta_b200.lowering import lower_game_config  # noqa: E402
from metta_b200.maps import arena_maps  # noqa: E402
from metta_b200.sharding import shard_seeds  # noqa: E402
from metta_b200.vecenv import GpuMettaGridVecEnv  # noqa: E402


class TokenPolicy(torch.nn.Module):
    """Bag of tokens: per agent, how often each feature id occurs and the sum of its values (two
    scatter-adds over the live tokens), then an MLP. Cheap enough that the env step still shows."""

    def __init__(self, n_actions: int, hidden: int = 256):
        super().__init__()
        self.body = torch.nn.Sequential(torch.nn.Linear(512, hidden), torch.nn.ReLU(), torch.nn.Linear(hidden, hidden),
                                        torch.nn.ReLU(), torch.nn.Linear(hidden, n_actions))

    def forward(self, obs: torch.Tensor, max_tokens: int) -> torch.Tensor:
        tok = obs[:, :max_tokens]                         # padding beyond the longest row carries nothing
        feat = tok[:, :, 1].long()                        # 255 = padding: lands in a column the MLP may ignore
        bag = torch.zeros((obs.shape[0], 512), device=obs.device, dtype=torch.float32)
        bag.scatter_add_(1, feat, torch.ones_like(feat, dtype=torch.float32))
        bag.scatter_add_(1, feat + 256, tok[:, :, 2].float() * (1.0 / 255.0))
        return self.body(bag.to(torch.bfloat16))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=4096, help="envs per GPU")
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--max-tokens", type=int, default=128, help="tokens the policy looks at (arena rows stay below ~105)")
    args = ap.parse_args()
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    env = GpuMettaGridVecEnv(cfg, args.envs, map_codes=arena_maps(low, 256, seed=1234), device=local,
                             seed=int(shard_seeds(rank, world, args.envs, 42)[0]))
    env.collect_infos = False
    policy = TokenPolicy(env.single_action_space.n).cuda().to(torch.bfloat16)
    obs, _ = env.reset()
    with torch.no_grad():
        for timed in (False, True):
            torch.cuda.synchronize()
            if world > 1:
                torch.distributed.barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps if timed else 20):
                logits = policy(obs, args.max_tokens)
                actions = torch.distributions.Categorical(logits=logits.float()).sample().to(torch.int32)
                obs, rewards, terminals, truncations, _ = env.step(actions)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
    rate = torch.tensor([args.envs * 24 * args.steps / dt], device="cuda", dtype=torch.float64)
    stats = env.reduce_episode_stats() if world > 1 else None
    if world > 1:
        torch.distributed.all_reduce(rate)
    if rank == 0:
        print(f"rollout with policy in the loop: {rate.item() / 1e6:.1f} M agent-steps/s on {world} GPU(s), "
              f"{args.envs} envs x 24 agents per GPU, {1e3 * dt / args.steps:.3f} ms per step (env + policy + sampling)")
    env.close()
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
