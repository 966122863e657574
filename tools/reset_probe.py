"""Cycle counts of the auto-reset path inside the step kernel (library built with -DMG_TIMING): every env of the
batch ends its episode on the same step. usage: MG_LIB_PATH=build_variants/lib_timing.so python tools/reset_probe.py"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.getcwd())
from metta_b200.batch import Batch
from metta_b200.builders import make_arena
from metta_b200.lowering import lower_game_config
from metta_b200.maps import arena_mapgen_pool, build_pool
N = int(sys.argv[1]) if len(sys.argv) > 1 else 512
A, T = 24, 200
cfg = make_arena(A); low = lower_game_config(cfg.game, 62, 62)
b = Batch(low, build_pool(low, arena_mapgen_pool(low)), N)
dev = 'cuda'
obs = torch.zeros((N*A, T, 3), dtype=torch.uint8, device=dev); term = torch.zeros(N*A, dtype=torch.uint8, device=dev)
trunc = torch.zeros_like(term); rew = torch.zeros(N*A, dtype=torch.float32, device=dev); act = torch.zeros(N*A, dtype=torch.int32, device=dev)
b.set_buffers(obs.data_ptr(), term.data_ptr(), trunc.data_ptr(), rew.data_ptr(), act.data_ptr())
b.reset((np.arange(N)+42).astype(np.uint32), (np.arange(N) % 256).astype(np.int32))
b.set_early_reset(np.full(N, 300, np.uint32))
flush = torch.empty(256*1024*1024, dtype=torch.uint8, device=dev)
for s in range(301):
    act.copy_(torch.randint(0, 5, (N*A,), dtype=torch.int32, device=dev))
    if s == 300: flush.fill_(1)
    b.step()
torch.cuda.synchronize()
sp = b.read_buffer(16, np.uint32, (N, 16)).astype(np.float64)
assert (b.read_env()["episodes"] == 1).all()
for name, i in (("load: loads+fold (contains the reset)", 0), ("  accumulate the finished episode", 12), ("  reset up to the assembler records", 13),
                ("  assembler records", 14), ("  count pass + folds", 15)):
    c = sp[:, i]; print('%-40s mean %8.0f p50 %8.0f max %8.0f' % (name, c.mean(), np.percentile(c, 50), c.max()))
