"""One step of the bench's steady state under a profiler: the batch is staggered over a whole episode length first
(profiling off), then cudaProfilerStart/Stop bracket one flushed step launched as two whole-batch kernels
(mg_profile_step) and one step launched the normal way (the sliced graph).
usage: ncu --profile-from-start off ... python tools/profile_steady.py [envs]"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.getcwd())
from metta_b200.batch import Batch
from metta_b200.builders import make_arena
from metta_b200.lowering import lower_game_config
from metta_b200.maps import arena_mapgen_pool, build_pool
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
A, T = 24, 200
cfg = make_arena(A); low = lower_game_config(cfg.game, 62, 62)
b = Batch(low, build_pool(low, arena_mapgen_pool(low)), N)
dev = 'cuda'
obs = torch.zeros((N*A, T, 3), dtype=torch.uint8, device=dev); term = torch.zeros(N*A, dtype=torch.uint8, device=dev)
trunc = torch.zeros_like(term); rew = torch.zeros(N*A, dtype=torch.float32, device=dev); act = torch.zeros(N*A, dtype=torch.int32, device=dev)
b.set_buffers(obs.data_ptr(), term.data_ptr(), trunc.data_ptr(), rew.data_ptr(), act.data_ptr())
b.reset((np.arange(N)+42).astype(np.uint32), (np.arange(N) % 256).astype(np.int32))
g = np.arange(N, dtype=np.int64)
b.set_early_reset((1 + (g * 6181) % 9999).astype(np.uint32))
for s in range(10000):
    act.copy_(torch.randint(0, 5, (N*A,), dtype=torch.int32, device=dev)); b.step()
flush = torch.empty(256*1024*1024, dtype=torch.uint8, device=dev)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for kind in ("kernels", "graph"):
    act.copy_(torch.randint(0, 5, (N*A,), dtype=torch.int32, device=dev)); flush.fill_(1); torch.cuda.synchronize()
    if kind == "kernels":
        print("profile_step ms", b.profile_step(torch.cuda.current_stream().cuda_stream))
    else:
        b.step(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
torch.cuda.profiler.stop()
b.close()
