"""BASELINE.json configs[4] in small: a rollout loop with a torch policy in it.

    python examples/rollout.py [--envs 4096] [--steps 200] [--shim kernel|torch]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/rollout.py

The env writes observations straight into the tensor the policy reads (no host copy, no extra
device copy), the policy's sampled actions go back as a device tensor, and everything is ordered on
one CUDA stream. The policy is a stand-in of the reference's token pipeline (strip 0xFF padding,
embed (feature, value) per token, pool, act: agent/src/metta/agent/components/obs_shim.py:166-196),
not a trained model. `--shim torch` recomputes the padding mask, the per-row flip point and their
maximum from the observations, as the reference's shim does for every batch; `--shim kernel` (default)
takes them from the observation kernel (mg_set_obs_aux: `env.obs_mask`, `env.row_lengths`,
`env.max_flip`). One process per GPU; envs are sharded by rank (metta_b200/sharding.py) and the only
collective is the all-reduce of episode statistics at the end.
"""

import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from metta_b200.builders import make_arena  # noqa: E402
from metta_b200.lowering import lower_game_config  # noqa: E402
from metta_b200.maps import arena_mapgen_pool  # noqa: E402
from metta_b200.sharding import shard_seeds  # noqa: E402
from metta_b200.vecenv import GpuMettaGridVecEnv  # noqa: E402


class TokenPolicy(torch.nn.Module):
    """Bag of tokens: per agent, how often each feature id occurs and the sum of its values (two
    scatter-adds over the live tokens), then an MLP. Cheap enough that the env step still shows."""

    def __init__(self, n_actions: int, hidden: int = 256):
        super().__init__()
        self.body = torch.nn.Sequential(torch.nn.Linear(512, hidden), torch.nn.ReLU(), torch.nn.Linear(hidden, hidden),
                                        torch.nn.ReLU(), torch.nn.Linear(hidden, n_actions))

    def forward(self, tok: torch.Tensor, pad: torch.Tensor) -> torch.Tensor:
        """tok u8 [B, L, 3] (already cut to the longest row), pad bool [B, L] (True = padding)."""
        feat = tok[:, :, 1].long()
        live = (~pad).float()
        bag = torch.zeros((tok.shape[0], 512), device=tok.device, dtype=torch.float32)
        bag.scatter_add_(1, feat, live)
        bag.scatter_add_(1, feat + 256, tok[:, :, 2].float() * (1.0 / 255.0) * live)
        return self.body(bag.to(torch.bfloat16))


def shim_torch(obs: torch.Tensor):
    """The reference's ObsShimTokens.forward, obs_shim.py:181-196, on the raw observation tensor."""
    M = obs.shape[1]
    coords = obs[..., 0]
    obs_mask = coords == 255
    flip_pts = obs_mask.int().argmax(dim=1)
    has_padding = obs_mask.any(dim=1)
    row_lengths = torch.where(has_padding, flip_pts, torch.full_like(flip_pts, M))
    max_flip = int(row_lengths.max())
    return obs[:, :max_flip], obs_mask[:, :max_flip]


def shim_kernel(env: GpuMettaGridVecEnv, obs: torch.Tensor):
    """The same two tensors from what the observation kernel already knows."""
    max_flip = int(env.max_flip)
    return obs[:, :max_flip], env.obs_mask[:, :max_flip]


def run_rollout(envs: int = 4096, steps: int = 200, warmup: int = 20, shim: str = "kernel", stagger: bool = False, quiet: bool = False) -> dict:
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    own_group = False
    if world > 1 and not torch.distributed.is_initialized():
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
        own_group = True
    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    env = GpuMettaGridVecEnv(cfg, envs, map_codes=arena_mapgen_pool(low), device=local,
                             seed=int(shard_seeds(rank, world, envs, 42)[0]), obs_mask=True)
    env.collect_infos = False
    policy = TokenPolicy(env.single_action_space.n).cuda().to(torch.bfloat16)
    obs, _ = env.reset()
    results = {}
    with torch.no_grad():
        for which in (("torch", "kernel") if shim == "both" else (shim,)):
            for timed in (False, True):
                torch.cuda.synchronize()
                if world > 1:
                    torch.distributed.barrier()
                t0 = time.perf_counter()
                for _ in range(steps if timed else warmup):
                    tok, pad = shim_kernel(env, obs) if which == "kernel" else shim_torch(obs)
                    logits = policy(tok, pad)
                    actions = torch.distributions.Categorical(logits=logits.float()).sample().to(torch.int32)
                    obs, rewards, terminals, truncations, _ = env.step(actions)
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
            results[which] = dt / steps
    best = "kernel" if "kernel" in results else shim
    rate = torch.tensor([envs * 24 / results[best]], device="cuda", dtype=torch.float64)
    if world > 1:
        env.reduce_episode_stats()
        torch.distributed.all_reduce(rate)
    out = {"agent_steps_per_s": float(rate), "ms_per_step": 1e3 * results[best], "shim": best,
           "ms_per_step_by_shim": {k: 1e3 * v for k, v in results.items()}, "world": world}
    if rank == 0 and not quiet:
        print(f"rollout with policy in the loop: {out['agent_steps_per_s'] / 1e6:.1f} M agent-steps/s on {world} GPU(s), "
              f"{envs} envs x 24 agents per GPU, ms per step (env + policy + sampling): "
              + ", ".join(f"{k} shim {v:.3f}" for k, v in out["ms_per_step_by_shim"].items()))
    env.close()
    if own_group:
        torch.distributed.destroy_process_group()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=4096, help="envs per GPU")
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--shim", default="both", choices=["kernel", "torch", "both"])
    args = ap.parse_args()
    run_rollout(args.envs, args.steps, shim=args.shim)


if __name__ == "__main__":
    main()
