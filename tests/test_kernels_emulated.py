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


@pytest.mark.parametrize("name,steps", [("kitchen_sink", 120), ("tiny_truncating", 17), ("clipped_start", 80)])
def test_emulated_kernels_match_oracle(name, steps, emu_lib):
    cfg, grid = SCENARIOS[name]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    ls = Lockstep(cfg, names_to_codes(low, grid), num_envs=3, seeds=[3, 4, 5], clib=emu_lib)
    ls.compare_outputs()
    rng = np.random.RandomState(11)
    for s in range(steps):
        ls.step(random_actions(rng, cfg, ls.lowered, ls.N * ls.A, invalid_every=29, step=s))
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
