"""Standalone time of the two kernels in the bench's steady state (mg_profile_step, L2 flushed), for A/B runs of
build variants: MG_LIB_PATH=build_variants/lib_x.so python tools/obs_ab.py"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.getcwd())
import bench
cfg, low, codes = bench._arena_inputs(256)
from metta_b200.maps import build_pool
dev = torch.device("cuda", 0); torch.cuda.set_device(0)
N = 4096
rig = bench.Rig(torch, low, build_pool(low, codes), N, 24, 200, 5, dev, (np.arange(N) + 42).astype(np.uint32), (np.arange(N) % 256).astype(np.int32), 0)
rig.stagger(10000, 0)
flush = bench.make_flush(torch, dev, "write+read")
a, o = [], []
for i in range(60):
    rig.fresh_actions(rig.act); flush(i); torch.cuda.synchronize()
    x, y = rig.batch.profile_step(rig.stream); a.append(x); o.append(y)
print("%-40s step kernel %.4f ms  obs kernel %.4f ms (median of 60; min obs %.4f)" % (os.environ.get("MG_LIB_PATH", "default"), np.median(a), np.median(o), np.min(o)))
