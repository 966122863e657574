"""When does the observation work start in hand-over mode (MG_STEP_MODE=flags, a -DMG_TIMING build)?
Per env: when its step started and ended, when the env's first observation CTA was dispatched, when it found
its env, and when its row was done. usage: MG_LIB_PATH=build_variants/lib_timing.so MG_STEP_MODE=flags python tools/handover_probe.py"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.getcwd())
from metta_b200.batch import Batch
from metta_b200.builders import make_arena
from metta_b200.lowering import lower_game_config
from metta_b200.maps import arena_mapgen_pool, build_pool
N, A, T = 4096, 24, 200
cfg = make_arena(A); low = lower_game_config(cfg.game, 62, 62)
b = Batch(low, build_pool(low, arena_mapgen_pool(low)), N)
dev = 'cuda'
obs = torch.zeros((N*A, T, 3), dtype=torch.uint8, device=dev); term = torch.zeros(N*A, dtype=torch.uint8, device=dev)
trunc = torch.zeros_like(term); rew = torch.zeros(N*A, dtype=torch.float32, device=dev); act = torch.zeros(N*A, dtype=torch.int32, device=dev)
b.set_buffers(obs.data_ptr(), term.data_ptr(), trunc.data_ptr(), rew.data_ptr(), act.data_ptr())
b.reset((np.arange(N)+42).astype(np.uint32), (np.arange(N) % 256).astype(np.int32))
g = np.arange(N, dtype=np.int64)
b.set_early_reset((1 + (g * 6181) % 9999).astype(np.uint32))
for s in range(3000):
    act.copy_(torch.randint(0, 5, (N*A,), dtype=torch.int32, device=dev)); b.step()
flush = torch.empty(256*1024*1024, dtype=torch.uint8, device=dev)
for s in range(5):
    act.copy_(torch.randint(0, 5, (N*A,), dtype=torch.int32, device=dev)); flush.fill_(s); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); b.step(torch.cuda.current_stream().cuda_stream); e1.record(); torch.cuda.synchronize()
print('mode', os.environ.get('MG_STEP_MODE', 'sliced'), 'step ms (CUDA events around the call)', e0.elapsed_time(e1))
sp = b.read_buffer(16, np.uint32, (N, 16)).astype(np.int64)
start = sp[:, 10]; cyc = sp[:, :10].sum(1) + sp[:, 11]
t0 = start.min()
rel = lambda x: ((x - t0) % 2**32) / 1000.0
s_start, s_end = rel(start), rel(start) + cyc / 1.92 / 1000.0
disp, go, done = rel(sp[:, 13]), rel(sp[:, 14]), rel(sp[:, 15])
pct = lambda x: ' '.join('%6.1f' % np.percentile(x, q) for q in (1, 10, 50, 90, 99, 100))
print('us after the first step warp started, percentiles 1 10 50 90 99 100 over envs')
print('step start                 ', pct(s_start))
print('step end                   ', pct(s_end))
print('obs CTA (group 0) dispatched', pct(disp))
print('obs CTA found its env       ', pct(go))
print('obs row done                ', pct(done))
print('wait inside the CTA         ', pct(go - disp))
print('env done -> obs starts      ', pct(go - s_end))
print('first step warp -> last group-0 row done: %.1f us' % done.max())
b.close()
