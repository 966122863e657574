"""Host-side logic and the C-ABI surface (no GPU needed)."""

import os
import re

import numpy as np
import pytest

from metta_b200 import capi
from metta_b200.builders import make_arena
from metta_b200.config import MettaGridConfig
from metta_b200.layout import L
from metta_b200.lowering import compute_features, lower_game_config, map_order, num_inventory_tokens
from metta_b200.maps import arena_maps, build_pool, codes_to_names, names_to_codes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(product_lib):
    header = open(os.path.join(ROOT, "include", "mettagrid_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(mg_\w+)\s*\(", header))
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    for name in declared:
        assert hasattr(product_lib, name), name
    assert product_lib.mg_abi_version() == L.MG_CONFIG_VERSION


def test_no_gpu_means_loud_error(product_lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from metta_b200.batch import Batch

    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    with pytest.raises(capi.MgError, match="no CUDA device"):
        Batch(low, build_pool(low, arena_maps(low, 1)), 4)
    from metta_b200.vecenv import GpuMettaGridVecEnv

    with pytest.raises(RuntimeError, match="no CPU simulator"):
        GpuMettaGridVecEnv(cfg, 4, map_codes=arena_maps(low, 1))


def test_unordered_map_orders(product_lib):
    # set_inventory({"heart":1,"ore_red":2,"laser":3}) iterates 8,0,6 -> 6,0,8
    assert [k for k, _ in map_order([(8, 1), (0, 2), (6, 3)])] == [6, 0, 8]
    assert [k for k, _ in map_order([(3, 1), (1, 2)])] == [1, 3]
    assert map_order([]) == [] and map_order([(5, 9)]) == [(5, 9)]


def test_arena_lowering_ids(product_lib):
    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    assert low.action_names == ["noop", "move_north", "move_south", "move_west", "move_east"]
    w = low.words
    assert w[L.MGC_NUM_ACTIONS] == 5 and w[L.MGC_NUM_RESOURCES] == 10 and w[L.MGC_NUM_TOKENS] == 200
    # type ids: 1 + rank of the object key in sorted order, agents 0 (mettagrid_c_config.py:54-55)
    assert low.type_names == {0: "agent", 1: "armory", 2: "assembler", 3: "generator_red", 4: "lasery", 5: "mine_red", 6: "wall"}
    # tag ids: rank in the sorted set of all tags (:80-96)
    assert low.tag_names == sorted(["agent", "armory", "assembler", "generator_red", "lasery", "mine_red", "wall"])
    feats = compute_features(cfg.game)
    assert len(feats) == 52 and feats[0][1] == "agent:group" and feats[12][1] == "inv:ore_red" and feats[13][1] == "inv:ore_red:p1"
    assert num_inventory_tokens(256) == 2 and num_inventory_tokens(100) == 3 and num_inventory_tokens(10) == 5
    for alias in ("agent.agent", "agent.default", "agent.red", "agent.team_0"):
        assert low.cell_codes[alias] == low.cell_codes["agent.agent"]


def test_config_errors_match_reference_messages(product_lib):
    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    with pytest.raises(RuntimeError, match="Unknown object type: lava"):
        names_to_codes(low, [["wall", "lava"]])
    bad = make_arena(24)
    bad.game.obs.width = 17
    with pytest.raises(ValueError, match="exceeds maximum packable size"):
        lower_game_config(bad.game, 62, 62)
    bad = make_arena(24)
    bad.game.actions.move.consumed_resources = {"mana": 1}
    with pytest.raises(ValueError, match="not in resource_names"):
        lower_game_config(bad.game, 62, 62)
    with pytest.raises(Exception):
        MettaGridConfig(game={"no_such_field": 1})


def test_map_pool_layout(product_lib):
    cfg = make_arena(24)
    low = lower_game_config(cfg.game, 62, 62)
    codes = arena_maps(low, 3, seed=4)
    pool = build_pool(low, codes)
    assert pool.grids.shape == (3, 62 * 62) and pool.agents.shape == (3, 24, 3)
    g = pool.grids.reshape(3, 62, 62)
    for m in range(3):
        # agents are numbered in row-major order of appearance (mettagrid_c.cpp:181-217)
        cells = np.argwhere(g[m] >> 14 == 1)
        idx = g[m][cells[:, 0], cells[:, 1]] & 0x3FFF
        assert np.array_equal(idx, np.arange(24))
        assert np.array_equal(pool.agents[m, :, :2], cells)
        assert pool.n_asm[m] == (g[m] >> 14 == 2).sum() == 4 * 22
    assert names_to_codes(low, codes_to_names(low, codes[0])).tolist() == codes[0].tolist()
    with pytest.raises(RuntimeError, match="agents"):
        build_pool(low, np.zeros((1, 62, 62), np.uint8))


def test_reference_config_dump_loads(product_lib):
    import json

    with open(os.path.join(ROOT, "tests", "golden", "arena.json")) as f:
        cfg = MettaGridConfig.from_reference(json.load(f))
    a = lower_game_config(cfg.game, 62, 62).words
    b = lower_game_config(make_arena(24).game, 62, 62).words
    assert np.array_equal(a, b), "the reference's own make_arena dump lowers to the same blob as builders.make_arena"
