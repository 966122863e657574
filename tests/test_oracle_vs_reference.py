"""Pins the oracle against the reference core itself (oracle/_ref, compiled from /root/reference).
Runs wherever the compiled reference travelled to; the part that needs the reference's Python
package only runs in the build container."""

import os

import numpy as np
import pytest
from harness import random_actions
from scenarios import SCENARIOS

from metta_b200.builders import make_arena
from metta_b200.config import MettaGridConfig
from metta_b200.lowering import lower_game_config
from metta_b200.maps import arena_maps, codes_to_names, names_to_codes
from oracle import ref_loader
from oracle.oracle import OracleEnv

needs_ref = pytest.mark.skipif(not ref_loader.available(), reason="oracle/_ref not built")


def lockstep(cfg, codes, low, seed, steps, invalid_every=0):
    from oracle.ref_bridge import RefEnv

    ref = RefEnv(cfg.game, codes_to_names(low, codes), seed)
    orc = OracleEnv(low.words, codes, seed)
    assert np.array_equal(ref.observations, orc.observations)
    rng = np.random.RandomState(seed)
    for s in range(steps):
        a = random_actions(rng, cfg, low, cfg.game.num_agents, invalid_every=invalid_every, step=s)
        ref.step(a)
        orc.step(a)
        assert np.array_equal(ref.observations, orc.observations), f"seed {seed} step {s}: observations"
        assert np.array_equal(ref.rewards, orc.rewards), f"seed {seed} step {s}: rewards"
        assert np.array_equal(ref.action_success(), orc.action_success()), f"seed {seed} step {s}: success"
        assert np.array_equal(ref.terminals, orc.terminals) and np.array_equal(ref.truncations, orc.truncations)
    assert np.array_equal(ref.episode_rewards(), orc.episode_rewards())
    return ref, orc


@needs_ref
@pytest.mark.parametrize("name,steps", [("kitchen_sink", 400), ("tiny_truncating", 12), ("clipped_start", 200), ("collective_rewards", 300)])
@pytest.mark.parametrize("seed", [1, 99])
def test_scenarios(name, steps, seed, product_lib):
    cfg, grid = SCENARIOS[name]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    ref, orc = lockstep(cfg, names_to_codes(low, grid), low, seed, steps, invalid_every=37)
    # stats by name: every stat the reference reports must equal the oracle's dense slot
    from metta_b200.stats import agent_stats_dict, game_stats_dict

    st = ref.episode_stats()
    game = game_stats_dict(low, orc.game_stats())
    for k, v in st["game"].items():
        assert game.get(k, 0.0) == pytest.approx(v), k
    astats = orc.agent_stats()
    from metta_b200.layout import L
    from metta_b200.maps import build_pool

    pool = build_pool(low, names_to_codes(low, grid))  # agent i = i-th agent cell in row-major order
    for i, ref_stats in enumerate(st["agent"]):
        cell = int(pool.agents[0, i, 2])
        grp = int(low.words[low.words[L.MGC_OFF_CELLTYPES] + cell * L.MG_CELL_WORDS + L.MG_CELL_PARAM])
        mine = agent_stats_dict(low, grp, astats[i], orc.invalid_ids()[i])
        for k, v in ref_stats.items():
            assert mine.get(k, 0.0) == pytest.approx(v), (i, k)  # zero-valued keys are omitted


@needs_ref
def test_arena_synthetic_maps(product_lib):
    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    codes = arena_maps(low, 2, seed=3)
    for m in range(2):
        lockstep(cfg, codes[m], low, 42 + m, 300)


@needs_ref
@pytest.mark.skipif(not os.path.isdir("/root/reference/packages/mettagrid/python/src"), reason="reference python tree absent")
def test_bridge_equals_reference_lowering(product_lib):
    """oracle/ref_bridge.py must configure the core exactly like the reference's own
    convert_to_cpp_game_config: same observations from both on arena."""
    ref_loader.load_reference_python()
    from mettagrid.builder.envs import make_arena as ref_make_arena
    from mettagrid.config.mettagrid_c_config import convert_to_cpp_game_config

    from oracle.ref_bridge import RefEnv

    ref_cfg = ref_make_arena(24)
    ref_cfg.game.map_builder.seed = 77
    grid = ref_cfg.game.map_builder.create().build_for_num_agents(24).grid
    cfg = MettaGridConfig.from_reference(ref_cfg)
    own = RefEnv(cfg.game, grid.tolist(), 5, c_cfg=convert_to_cpp_game_config(ref_cfg.game.model_dump()))
    bridged = RefEnv(cfg.game, grid.tolist(), 5)
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    assert np.array_equal(low.words, lower_game_config(make_arena(24).game, 62, 62).words)
    rng = np.random.RandomState(0)
    for s in range(200):
        a = rng.randint(0, 5, size=24).astype(np.int32)
        own.step(a)
        bridged.step(a)
        assert np.array_equal(own.observations, bridged.observations) and np.array_equal(own.rewards, bridged.rewards)


@needs_ref
@pytest.mark.parametrize("name,seeds,steps", [("dense_conflict", 64, 2000), ("crowd_48", 12, 1500)])
def test_crowded_rooms_many_seeds(name, seeds, steps, product_lib):
    """Rooms where nearly every action meets another agent (attack, transfer, swap with frozen agents,
    shared assembler neighbourhoods, chests within reach of an assembler, loot that removes a limit
    modifier): the oracle against the compiled reference on many seeds, every byte of every step."""
    cfg, grid = SCENARIOS[name]()
    low = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    codes = names_to_codes(low, grid)
    for seed in range(300, 300 + seeds):
        lockstep(cfg, codes, low, seed, steps, invalid_every=41)


@needs_ref
def test_arena_whole_episode(product_lib):
    """A whole arena episode (10 000 steps, MapGen map): inventories at their limits, float stat totals
    beyond 2^24, the generator regenerated 190 times."""
    from metta_b200.maps import arena_mapgen_pool

    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    codes = arena_mapgen_pool(low, 2)[1]
    from oracle.ref_bridge import RefEnv

    ref = RefEnv(cfg.game, codes_to_names(low, codes), 43)
    orc = OracleEnv(low.words, codes, 43)
    rng = np.random.RandomState(1)
    for s in range(10000):
        a = rng.randint(0, 5, size=24).astype(np.int32)
        ref.step(a)
        orc.step(a)
        assert np.array_equal(ref.observations, orc.observations) and np.array_equal(ref.rewards, orc.rewards), f"step {s}"
    assert ref.truncations.all() or ref.terminals.all()
    assert np.array_equal(ref.terminals, orc.terminals) and np.array_equal(ref.truncations, orc.truncations)
    assert np.array_equal(ref.episode_rewards(), orc.episode_rewards())


@needs_ref
@pytest.mark.skipif(not os.path.isdir(os.path.join(os.path.dirname(ref_loader.__file__), "_ref_ordered")), reason="oracle/_ref_ordered not built")
def test_clipper_picks_against_the_id_ordered_reference_build(product_lib):
    """Clipper::pick_assembler_to_clip walks std::set<Assembler*> in heap-address order (clipper.hpp:207-229),
    which no two processes share. oracle/build_ref_ordered.sh builds the same reference sources with those
    sets ordered by object id (a force-included comparator, nothing else changed); against that build the
    oracle must agree on every clip: several seeds, about 400 clip/unclip events each."""
    import subprocess
    import sys

    script = os.path.join(os.path.dirname(os.path.abspath(__file__)), "clipper_ordered_check.py")
    env = dict(os.environ, MG_REF_DIR="_ref_ordered")
    out = subprocess.run([sys.executable, script], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]
