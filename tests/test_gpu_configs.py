"""BASELINE.json configs[2] and configs[3] at their full sizes on one GPU: a sampled lockstep against
the oracle (with MettaGridPufferEnv's auto-reset rule restated for the oracle) plus properties that
hold for every env of the batch."""

import numpy as np
import pytest
from golden_util import load
from harness import HostBuffers, random_actions

from metta_b200.batch import Batch
from metta_b200.builders import make_navigation
from metta_b200.lowering import lower_game_config
from metta_b200.maps import build_pool, names_to_codes, random_maps

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(product_lib):
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return product_lib


def _run(cfg, low, codes, n, steps, sample, map_stride, twin_offset, seed0=100):
    """Steps n envs; env i and env i + twin_offset get the same seed, map and actions, so their
    outputs must agree (a whole-batch property); envs in `sample` are checked against the oracle."""
    from oracle.oracle import OracleEnv

    A, T = cfg.game.num_agents, cfg.game.obs.num_tokens
    M = len(codes)
    b = Batch(low, build_pool(low, codes), n)
    b.set_options(auto_reset=True, map_stride=map_stride)
    buf = HostBuffers(n * A, T, cuda=True)
    b.set_buffers(*buf.ptrs)
    base = np.arange(n) % twin_offset
    seeds = (base + seed0).astype(np.uint32)
    maps = (base % M).astype(np.int32)
    b.reset(seeds, maps)
    cur_map = {i: int(maps[i]) for i in sample}
    orcs = {i: OracleEnv(low.words, codes[cur_map[i]], int(seeds[i])) for i in sample}
    rng = np.random.RandomState(1)
    episodes = 0
    for s in range(steps):
        a = random_actions(rng, cfg, low, twin_offset * A).reshape(twin_offset, A)
        a = np.concatenate([a] * (n // twin_offset)).reshape(-1)
        buf.set_actions(a)
        b.step()
        an = a.reshape(n, A)
        obs, rew, term, trunc = buf.fetch()
        for i in sample:
            o = orcs[i]
            if o.terminals.all() or o.truncations.all():  # mettagrid_puffer_env.py:158-164: new sim, same seed
                cur_map[i] = (cur_map[i] + map_stride) % M
                o.close()
                o = orcs[i] = OracleEnv(low.words, codes[cur_map[i]], int(seeds[i]))
                episodes += 1
            o.step(an[i])
            sl = slice(i * A, (i + 1) * A)
            assert np.array_equal(obs[sl], o.observations), f"step {s} env {i}"
            assert np.array_equal(rew[sl], o.rewards) and np.array_equal(trunc[sl].astype(bool), o.truncations)
    # every env against its twin
    o3 = obs.reshape(n // twin_offset, twin_offset * A, T, 3)
    assert all(np.array_equal(o3[0], o3[k]) for k in range(1, n // twin_offset)), "twins (same seed, map, actions) diverged"
    # structure of every row: tokens form a prefix, global token first
    flat = obs.reshape(n * A, T, 3)
    empty = (flat == 255).all(axis=2)
    assert np.all(np.diff(empty.astype(np.int8), axis=1) >= 0)
    assert np.all(flat[:, 0, 1] == low.feature_ids["episode_completion_pct"])
    grid = b.read_grid()
    ag = b.read_agents()
    assert np.all(((grid >> 14) == 1).sum(axis=(1, 2)) == A)
    e, k = np.repeat(np.arange(n), A), np.tile(np.arange(A), n)
    assert np.array_equal(grid[e, ag["r"].reshape(-1), ag["c"].reshape(-1)], (0x4000 | k).astype(np.uint16))
    env = b.read_env()
    b.close()
    return env, episodes


def test_navigation_16k_envs_with_auto_reset(lib):
    """configs[2]: object-use maps (walls, heart assemblers), 16384 envs, episodes of 24 steps so that
    every env goes through the auto-reset from the HBM map pool several times."""
    cfg = make_navigation(num_agents=4, max_steps=24)
    cfg.game.episode_truncates = True
    low = lower_game_config(cfg.game, 40, 40)
    codes = random_maps(low, 64, 40, 40, {"wall": 120, "assembler": 24, "agent.agent": 4}, seed=3)
    env, episodes = _run(cfg, low, codes, 16384, 75, sample=[0, 1, 63, 64, 4097, 8191], map_stride=5, twin_offset=8192)
    assert np.all(env["episodes"] == 3) and np.all(env["current_step"] == 3) and episodes == 3 * 6


def test_cogames_mission_8k_envs(lib):
    """configs[3]: a cogames mission (the reference's own config and map from the golden fixture: 157
    actions, vibe-keyed assemblers, chests, chargers, regen, clipper with clip_period 0), 8192 envs."""
    cfg, grid, _ = load("cogames_vibe_check")
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    codes = names_to_codes(low, grid)[None]
    env, _ = _run(cfg, low, codes, 8192, 60, sample=[0, 7, 4095, 4096, 8191], map_stride=0, twin_offset=4096)
    assert np.all(env["current_step"] == 60)


def test_cogames_clipper_mission_lockstep(lib):
    """RepairMission (clipper with clip_period 50) on the GPU against the oracle, 48 envs x 300 steps. (Which
    assembler a clip picks is pinned against the reference by the clipper_spread / cogames_clip_period fixtures.)"""
    import json
    import os

    from golden_util import GOLDEN
    from harness import Lockstep

    from metta_b200.config import MettaGridConfig

    with open(os.path.join(GOLDEN, "cogames_repair_config.json")) as f:
        cfg = MettaGridConfig.from_reference(json.load(f))
    grid = np.load(os.path.join(GOLDEN, "cogames_repair_config.npz"))["grid"]
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    n = 48
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=n, seeds=np.arange(n) + 200, cuda=True)
    rng = np.random.RandomState(32)
    for s in range(300):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A))
    ls.compare_state()
    assert (ls.batch.read_assemblers()["flags"] & 1).any()
    ls.close()


def test_arena_long_run_with_episode_ends(lib):
    """Arena, 1024 envs, 2000 steps with episodes of 450 steps (so every env is rebuilt from the map pool
    four times, with map rotation): sampled envs in lockstep with the oracle the whole way."""
    from metta_b200.builders import make_arena
    from metta_b200.maps import arena_maps

    cfg = make_arena(24, max_steps=450)
    cfg.game.episode_truncates = True
    low = lower_game_config(cfg.game, 62, 62)
    codes = arena_maps(low, 32, seed=21)
    env, episodes = _run(cfg, low, codes, 1024, 2000, sample=[0, 511, 513, 1023], map_stride=3, twin_offset=512, seed0=900)
    assert np.all(env["episodes"] == 4) and episodes == 4 * 4
