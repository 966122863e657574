"""Times the observation kernel alone (mg_compute_observations) with CUDA events: back to back, and with an L2 flush before each launch."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.getcwd())
from metta_b200.batch import Batch
from metta_b200.builders import make_arena
from metta_b200.lowering import lower_game_config
from metta_b200.maps import arena_maps, build_pool
N, A, T = 4096, 24, 200
cfg = make_arena(A); low = lower_game_config(cfg.game, 62, 62)
pool = build_pool(low, arena_maps(low, 256, seed=1234))
b = Batch(low, pool, N)
dev = 'cuda'
obs = torch.zeros((N*A, T, 3), dtype=torch.uint8, device=dev); term = torch.zeros(N*A, dtype=torch.uint8, device=dev); trunc = torch.zeros_like(term)
rew = torch.zeros(N*A, dtype=torch.float32, device=dev); act = torch.randint(0, 5, (N*A,), dtype=torch.int32).to(dev)
b.set_buffers(obs.data_ptr(), term.data_ptr(), trunc.data_ptr(), rew.data_ptr(), act.data_ptr())
b.reset((np.arange(N)+42).astype(np.uint32), (np.arange(N) % 256).astype(np.int32))
for s in range(30):
    act.copy_(torch.randint(0, 5, (N*A,), dtype=torch.int32)); b.step()
torch.cuda.synchronize()
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(256*1024*1024, dtype=torch.uint8, device=dev)
def timed(n, do_flush):
    ts = []
    for i in range(n):
        if do_flush: flush.fill_(i)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); b.compute_observations(st); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return np.array(ts[3:])
w = timed(30, False); c = timed(30, True)
print('obs kernel alone: warm L2 mean %.1f us min %.1f | flushed L2 mean %.1f us min %.1f' % (w.mean()*1e3, w.min()*1e3, c.mean()*1e3, c.min()*1e3))
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(50): b.compute_observations(st)
e1.record(); torch.cuda.synchronize(); print('50 back to back: %.1f us each' % (e0.elapsed_time(e1)/50*1e3))
for do_flush in (False, True):
    a_l, o_l = [], []
    for i in range(20):
        act.copy_(torch.randint(0, 5, (N*A,), dtype=torch.int32).to(dev))
        if do_flush: flush.fill_(i)
        torch.cuda.synchronize()
        a_ms, o_ms = b.profile_step(st); a_l.append(a_ms); o_l.append(o_ms)
    print('profile_step flush=%s: step %.1f us obs %.1f us' % (do_flush, np.mean(a_l[3:])*1e3, np.mean(o_l[3:])*1e3))
# obs right after a step, timed separately
o2 = []
for i in range(20):
    act.copy_(torch.randint(0, 5, (N*A,), dtype=torch.int32).to(dev)); flush.fill_(i)
    b.step(st); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); b.compute_observations(st); e1.record(); torch.cuda.synchronize(); o2.append(e0.elapsed_time(e1))
print('obs recomputed after a full step (L2 holds the step): %.1f us' % (np.mean(o2[3:])*1e3))
