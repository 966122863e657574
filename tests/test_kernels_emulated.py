"""Executes the simulator's kernel SOURCES on the CPU (tests/simt_emu) in lockstep with the oracle.
This checks kernel logic where no GPU is available; the GPU parity tests (-m gpu) run the same
comparisons against the compiled sm_100a kernels."""

import numpy as np
import pytest
from harness import Lockstep, random_actions
from scenarios import SCENARIOS

from metta_b200.builders import make_arena
from metta_b200.maps import arena_maps, names_to_codes
from metta_b200.lowering import lower_game_config


@pytest.mark.parametrize("name,steps", [("kitchen_sink", 120), ("tiny_truncating", 17), ("clipped_start", 80), ("collective_rewards", 90)])
def test_emulated_kernels_match_oracle(name, steps, emu_lib):
    cfg, grid = SCENARIOS[name]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=3, seeds=[3, 4, 5], clib=emu_lib)
    ls.compare_outputs()
    rng = np.random.RandomState(11)
    for s in range(steps):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A, invalid_every=3 if name == "collective_rewards" else 29, step=s))
        if name != "tiny_truncating" and s % 40 == 39:
            ls.compare_state()
    ls.close()


@pytest.mark.parametrize("name,steps", [("cogames_vibe_check", 60), ("cogames_machina_chests", 25)])
def test_emulated_cogames_missions(name, steps, emu_lib):
    """cogames missions (157 actions, vibe-keyed protocols, chests, chargers, clip_period 0): the
    reference's own config and map from the golden fixture, kernels vs oracle in lockstep."""
    from golden_util import load

    cfg, grid, _ = load(name)
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=2, seeds=[5, 6], clib=emu_lib)
    ls.compare_outputs()
    rng = np.random.RandomState(21)
    for s in range(steps):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A))
    ls.compare_state()
    ls.close()


def test_emulated_cogames_clipper(emu_lib):
    """RepairMission: clipper with clip_period 50 (its own generator copy, infection weights, unclip
    protocols). Oracle <-> kernels only: the reference's pick order depends on heap addresses."""
    import json
    import os

    from golden_util import GOLDEN

    from metta_b200.config import MettaGridConfig

    with open(os.path.join(GOLDEN, "cogames_repair_config.json")) as f:
        cfg = MettaGridConfig.from_reference(json.load(f))
    grid = np.load(os.path.join(GOLDEN, "cogames_repair_config.npz"))["grid"]
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=2, seeds=[8, 9], clib=emu_lib)
    rng = np.random.RandomState(31)
    for s in range(160):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A))
    ls.compare_state()
    asm = ls.batch.read_assemblers()
    assert (asm["flags"] & 1).any(), "the clipper should have clipped something within 160 steps at period 50"
    ls.close()


def test_emulated_arena(emu_lib):
    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    ls = Lockstep(cfg, arena_maps(low, 2, seed=9), num_envs=2, seeds=[42, 43], clib=emu_lib)
    ls.compare_outputs()
    rng = np.random.RandomState(2)
    for s in range(60):
        ls.step(rng.randint(0, 5, size=ls.N * ls.A))
    ls.compare_state()
    ls.close()


def test_emulated_chunked_launch(emu_lib, monkeypatch):
    """The step is launched in slices of envs (MG_CHUNKS); slices must tile the batch exactly."""
    monkeypatch.setenv("MG_CHUNKS", "3")
    cfg, grid = SCENARIOS["tiny_truncating"]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=7, seeds=np.arange(7) + 30, clib=emu_lib)
    rng = np.random.RandomState(4)
    for s in range(14):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A))
    ls.close()


def test_emulated_obs_aux(emu_lib):
    """mg_set_obs_aux: row lengths, their maximum and the padding mask equal what the policy's first op
    computes from the observations (agent/src/metta/agent/components/obs_shim.py:181-196)."""
    cfg, grid = SCENARIOS["kitchen_sink"]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=3, seeds=[3, 4, 5], clib=emu_lib)
    NA, T = ls.N * ls.A, ls.T
    row_len = np.full(NA, -1, np.int32)
    max_flip = np.full(1, -1, np.int32)
    mask = np.full((NA, T), 7, np.uint8)
    ls.batch.set_obs_aux(row_len.ctypes.data, max_flip.ctypes.data, mask.ctypes.data)
    rng = np.random.RandomState(5)
    for s in range(30):
        ls.step(random_actions(rng, cfg, ls.lowered, NA))
        obs = ls.buf.obs
        pad = obs[:, :, 0] == 255
        expect = np.where(pad.any(1), pad.argmax(1), T)
        assert np.array_equal(row_len, expect)
        assert max_flip[0] == expect.max()
        assert np.array_equal(mask.astype(bool), pad)
    ls.close()


@pytest.mark.parametrize("name,envs,steps", [("dense_conflict", 3, 150), ("crowd_48", 2, 60)])
def test_emulated_crowded_rooms(name, envs, steps, emu_lib):
    """The order-independence classifier where conflicts dominate (and, with 48 agents, the 64-agent
    instantiation of the step kernel whose agents all take the ordered pass)."""
    cfg, grid = SCENARIOS[name]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=envs, seeds=np.arange(envs) + 70, clib=emu_lib)
    ls.compare_outputs()
    rng = np.random.RandomState(12)
    for s in range(steps):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A, invalid_every=23, step=s))
        if s % 50 == 49:
            ls.compare_state()
    ls.compare_state()
    ordered = ls.batch.read_env()["spare"][:, 0].sum() / float(envs * ls.A * steps)
    st = ls.batch.read_agent_stats().sum(axis=(0, 1))
    from metta_b200.layout import L
    assert st[L.MG_AS_SWAP] > 0 and st[L.MG_AS_TRANSFER_COUNT] > 0 and st[L.MG_AS_ATTACK_HIT : L.MG_AS_ATTACK_HIT + 8].sum() > 0
    assert ordered > 0.6, f"only {ordered:.2f} of the agents went through the ordered pass"
    ls.close()


def test_emulated_episode_accumulators(emu_lib):
    """The finished-episode accumulators (what reduce_episode_stats all-reduces) equal the sums of the per-env
    stats and episode rewards as they stood when the episodes ended."""
    cfg, grid = SCENARIOS["tiny_truncating"]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    n = 5
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=n, seeds=np.arange(n) + 3, clib=emu_lib)
    rng = np.random.RandomState(4)
    want = None
    for s in range(1, 12):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A))
        if s in (5, 10):  # episodes end here (max_steps 5); the next step folds them into the accumulators
            st = ls.batch.read_agent_stats().astype(np.float64).sum(axis=(0, 1))
            gs = ls.batch.read_game_stats().astype(np.float64).sum(axis=0)
            rw = ls.batch.read_episode_rewards().astype(np.float64).sum()
            cur = np.concatenate([[float(n), rw], st, gs])
            want = cur if want is None else want + cur
    acc = ls.batch.read_episode_accum()
    assert acc[0] == 2 * n
    assert np.allclose(acc, want, rtol=0, atol=1e-3), np.argwhere(~np.isclose(acc, want, rtol=0, atol=1e-3))[:5].tolist()
    ls.close()


def test_emulated_clipper_picks_fixture(emu_lib):
    """The kernels against the reference's own clipper code (fixture from the id-ordered reference build):
    which assembler gets clipped, with which unclip protocol, and everything downstream, 400 steps."""
    from golden_util import load

    cfg, grid, d = load("clipper_spread")
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=1, seeds=[int(d["seed"])], clib=emu_lib)
    assert np.array_equal(ls.buf.obs, d["obs"][0])
    for s in range(400):
        ls.step(d["actions"][s])
        assert np.array_equal(ls.buf.obs, d["obs"][s + 1]), f"step {s}"
        assert np.array_equal(ls.buf.rew, d["rewards"][s])
    ls.close()


def _same_grid_objects(ref: dict, mine: dict) -> None:
    assert set(ref) == set(mine)
    for k, ro in ref.items():
        mo = mine[k]
        assert set(ro) == set(mo), (k, set(ro) ^ set(mo))
        for f, v in ro.items():
            if f == "protocols":
                assert sorted(repr({**x, "inputs": dict(x["inputs"]), "outputs": dict(x["outputs"])}) for x in v) == sorted(map(repr, mo[f])), (k, f)
            elif f == "vibe_transfers":
                assert {a: dict(b) for a, b in v.items()} == mo[f], (k, f)
            elif isinstance(v, dict):
                assert list(v.items()) == list(mo[f].items()), (k, f, v, mo[f])  # iteration order included
            elif isinstance(v, float):
                assert abs(v - mo[f]) < 1e-6, (k, f)
            else:
                assert v == mo[f], (k, f, v, mo[f])


@pytest.mark.parametrize("name", ["kitchen_sink", "collective_rewards"])
def test_emulated_grid_objects_match_the_reference(name, emu_lib):
    """grid_objects() dicts built from the C-ABI records (metta_b200/grid_objects.py) against the compiled
    reference's own MettaGrid::grid_objects: ids, keys, feature entries, inventories in iteration order,
    assembler protocol views, chest transfers."""
    from oracle import ref_loader

    if not ref_loader.available():
        pytest.skip("oracle/_ref not built")
    from metta_b200.grid_objects import grid_objects

    cfg, grid = SCENARIOS[name]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=2, seeds=[3, 4], clib=emu_lib, checker="reference")
    rng = np.random.RandomState(1)
    for s in range(60):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A))
        if s % 12 == 0:
            for i in range(2):
                _same_grid_objects(ls.checkers[i].grid_objects(), grid_objects(ls.batch, i))
    sub = grid_objects(ls.batch, 0, 1, 4, 2, 7, ("wall",))
    ref = ls.checkers[0].sim.grid_objects(1, 4, 2, 7, ["wall"])
    _same_grid_objects(ref, sub)
    ls.close()


def test_emulated_replay_matches_the_reference_writer(emu_lib):
    from harness import check_replay_against_reference

    check_replay_against_reference(clib=emu_lib)
