import os, sys, time, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
os.environ["MG_STEP_MODE"] = "flags"
from harness import Lockstep
from metta_b200.builders import make_arena
from metta_b200.lowering import lower_game_config
from metta_b200.maps import arena_maps
cfg = make_arena(24, max_steps=25); cfg.game.episode_truncates = True
low = lower_game_config(cfg.game, 62, 62)
n = int(sys.argv[1])
ls = Lockstep(cfg, arena_maps(low, 4, seed=6), num_envs=n, seeds=np.arange(n) + 70, cuda=True)
rng = np.random.RandomState(9)
t0 = time.time()
for s in range(6):
    ls.step(rng.randint(0, 5, size=n * 24), compare=(s == 5))
ls.batch.sync()
print('flags mode', n, 'envs: 6 steps ok in %.1f s' % (time.time() - t0))
ls.close()
