"""Step / observation kernel split (mg_profile_step, L2 flushed) for bench.py's other_configs workloads.
usage: python tools/other_config_split.py"""
import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
from metta_b200.builders import make_navigation
from metta_b200.lowering import lower_game_config
from metta_b200.maps import build_pool, names_to_codes, random_maps
from golden_util import load

dev = torch.device("cuda", 0)
flush = bench.make_flush(torch, dev, "write+read")


def split(name, rig, warm):
    for _ in range(warm):
        rig.one_step()
    a, o = [], []
    for i in range(30):
        rig.fresh_actions(rig.act); flush(i); torch.cuda.synchronize()
        x, y = rig.batch.profile_step(rig.stream); a.append(x); o.append(y)
    ms = rig.timed_steps(40, flush)
    if "timing" in os.environ.get("MG_LIB_PATH", ""):  # a -DMG_TIMING build: per-phase cycle counts of the last step
        sp = rig.batch.read_buffer(16, np.uint32, (rig.N, 16)).astype(np.float64); sp[:, 10] = 0
        names = ['load: loads+fold', 'load: reset/stage', 'load: rest', 'decide: intents', 'decide: shuffle draws', 'decide: conflicts+apply',
                 'act parallel', 'act ordered', 'post', 'finish: rewards+records', '', 'finish: write-back']
        tot = sp[:, :12].sum(1)
        print('  per-env cycles: mean %.0f p50 %.0f p99 %.0f' % (tot.mean(), np.percentile(tot, 50), np.percentile(tot, 99)))
        for i, nm in enumerate(names):
            if nm: print('  %-28s mean %7.0f p99 %7.0f' % (nm, sp[:, i].mean(), np.percentile(sp[:, i], 99)))
    print(f"{name}: step kernel {np.mean(a)*1e3:.1f} us, observation kernel {np.mean(o)*1e3:.1f} us, whole step {np.mean(ms)*1e3:.1f} us "
          f"({rig.N} envs x {rig.A} agents, {rig.N * rig.A / np.mean(ms) / 1e3:.1f} M agent-steps/s)")
    rig.close()


cfg = make_navigation(num_agents=4, max_steps=100); cfg.game.episode_truncates = True
low = lower_game_config(cfg.game, 40, 40)
codes = random_maps(low, 64, 40, 40, {"wall": 120, "assembler": 24, "agent.agent": 4}, seed=3)
n = 16384
rig = bench.Rig(torch, low, build_pool(low, codes), n, 4, int(cfg.game.obs.num_tokens), len(low.action_names), dev,
                ((np.arange(n) + 7) & 0xFFFFFFFF).astype(np.uint32), (np.arange(n) % 64).astype(np.int32), 100)
rig.batch.set_options(auto_reset=True, map_stride=5); rig.stagger(100, 0)
split("navigation_16k", rig, 10)
cfg, grid, _ = load("cogames_vibe_check")
low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
n, A = 8192, cfg.game.num_agents
rig = bench.Rig(torch, low, build_pool(low, names_to_codes(low, grid)[None]), n, A, int(cfg.game.obs.num_tokens), len(low.action_names), dev,
                ((np.arange(n) + 11) & 0xFFFFFFFF).astype(np.uint32), np.zeros(n, np.int32), 200)
split("cogames_8k", rig, 160)
