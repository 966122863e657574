"""How the share of agents that need the ordered pass, the tokens per row and the assembler uses develop over an
arena episode (no staggering: all envs start together). usage: python tools/ordered_probe.py [envs]"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.getcwd())
from metta_b200.batch import Batch
from metta_b200.builders import make_arena
from metta_b200.lowering import lower_game_config
from metta_b200.maps import arena_mapgen_pool, build_pool
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
A, T = 24, 200
cfg = make_arena(A); low = lower_game_config(cfg.game, 62, 62)
b = Batch(low, build_pool(low, arena_mapgen_pool(low)), N)
dev = 'cuda'
obs = torch.zeros((N*A, T, 3), dtype=torch.uint8, device=dev); term = torch.zeros(N*A, dtype=torch.uint8, device=dev)
trunc = torch.zeros_like(term); rew = torch.zeros(N*A, dtype=torch.float32, device=dev)
act = torch.zeros(N*A, dtype=torch.int32, device=dev)
rl = torch.zeros(N*A, dtype=torch.int32, device=dev)
b.set_buffers(obs.data_ptr(), term.data_ptr(), trunc.data_ptr(), rew.data_ptr(), act.data_ptr())
b.set_obs_aux(rl.data_ptr(), 0, 0)
b.reset((np.arange(N)+42).astype(np.uint32), (np.arange(N) % 256).astype(np.int32))
prev_o = 0; prev_u = 0
for s in range(1, 10001):
    act.copy_(torch.randint(0, 5, (N*A,), dtype=torch.int32, device=dev)); b.step()
    if s in (50, 200, 500, 1000, 2000, 3000, 5000, 7000, 9999):
        torch.cuda.synchronize()
        o = int(b.read_env()["spare"][:, 0].astype(np.int64).sum()); u = int(b.read_assemblers()["uses_count"].astype(np.int64).sum())
        ag = b.read_agents()
        inv = ag["amt"].astype(np.int64).sum(axis=2)
        # neighbours: agents with another agent in the 8 cells around
        r, c = ag["r"].astype(int), ag["c"].astype(int)
        near = (np.maximum(np.abs(r[:, :, None] - r[:, None, :]), np.abs(c[:, :, None] - c[:, None, :])) == 1).any(2).mean()
        print(f"step {s:5d}: ordered agents/env/step so far {o / N / s:.3f}  uses/env/step so far {u / N / s:.4f}  tokens/row {rl.float().mean().item():.1f} max {rl.max().item()}  items/agent {inv.mean():.1f}  agents with an adjacent agent {near:.3f}")
