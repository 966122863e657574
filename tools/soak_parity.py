"""Long lockstep run on the GPU against the oracle (one-off check beside the test suite): arena, E envs,
S steps across an episode end (10 000 steps), observations/rewards/flags compared every step, the whole
state every 1000 steps. usage: python tools/soak_parity.py [envs] [steps]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from harness import Lockstep
from metta_b200.builders import make_arena
from metta_b200.lowering import lower_game_config
from metta_b200.maps import arena_mapgen_pool

E = int(sys.argv[1]) if len(sys.argv) > 1 else 48
S = int(sys.argv[2]) if len(sys.argv) > 2 else 10400
cfg = make_arena(24)
low = lower_game_config(cfg.game, 62, 62)
codes = arena_mapgen_pool(low)[:16]
ls = Lockstep(cfg, codes, num_envs=E, seeds=np.arange(E) + 900, cuda=True)
rng = np.random.RandomState(123)
t0 = time.time()
for s in range(S):
    ls.step(rng.randint(0, 5, size=E * 24))
    if (s + 1) % 1000 == 0:
        ls.compare_state()
        print(f"step {s + 1}: outputs and state equal ({time.time() - t0:.0f} s)", flush=True)
ls.compare_state()
print("ordered agents so far:", int(ls.batch.read_env()["spare"][:, 0].sum()) if "spare" in ls.batch.read_env().dtype.names else "n/a")
print(f"soak ok: {E} envs x {S} steps")
ls.close()
