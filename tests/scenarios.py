"""Shared test scenarios: configs (in the reference's schema) and ASCII maps.

Each scenario exercises a slice of the hot path; together they cover moves and
blocking, swap with frozen agents, vibe-triggered attack/transfer, assemblers (cooldown, partial
usage, max uses, multi-vibe protocols, chest search), chests, regen, damage, shared limits with
modifiers, consumed resources, goal/compass tokens, token truncation, 8-way moves and teams.
"""

from __future__ import annotations

import numpy as np

from metta_b200.config import MettaGridConfig

VIBES = ["default", "charger", "carbon", "weapon", "shield", "heart", "gear"]


def _vibes(names):
    return [{"symbol": "", "name": n, "category": "misc"} for n in names]


def ascii_map(rows: list[str], legend: dict[str, str]) -> np.ndarray:
    return np.array([[legend[ch] for ch in row] for row in rows], dtype="<U20")


def kitchen_sink() -> tuple[MettaGridConfig, np.ndarray]:
    resources = ["energy", "carbon", "oxygen", "heart", "weapon", "shield", "gear", "battery"]
    game = {
        "resource_names": resources,
        "num_agents": 8,
        "max_steps": 0,
        "obs": {"width": 7, "height": 9, "num_tokens": 48, "token_value_base": 100},
        "global_obs": {"compass": True, "goal_obs": True},
        "inventory_regen_interval": 3,
        "agents": [
            {
                "team_id": t,
                "tags": ["agent", f"team{t}"],
                "freeze_duration": 4,
                "initial_vibe": 0,
                "inventory": {
                    "default_limit": 30,
                    "limits": {
                        "energy": {"limit": 120, "resources": ["energy"]},
                        "cargo": {"limit": 20, "resources": ["carbon", "oxygen"], "modifiers": {"gear": 10}},
                        "heart": {"limit": 300, "resources": ["heart"]},
                    },
                    "initial": {"energy": 100, "gear": 1, "carbon": 4, "weapon": 3, "shield": 2, "heart": 150},
                    "regen_amounts": {"default": {"energy": 2}, "charger": {"energy": 5, "oxygen": 1}},
                },
                "rewards": {
                    "inventory": {"heart": 0.5, "carbon": 0.125},
                    "inventory_max": {"carbon": 1.0},
                    "stats": {"chest.heart.deposited_by_agent": 1.0, "action.move.success": 0.01, "assembler.heart.created": 0.25},
                    "stats_max": {"action.move.success": 0.3},
                },
                "diversity_tracked_resources": ["carbon", "oxygen", "heart"],
                "damage": {"threshold": {"battery": 3}, "resources": {"carbon": 1, "oxygen": 0, "shield": 0}},
            }
            for t in (0, 0, 0, 0, 1, 1, 1, 1)
        ],
        "actions": {
            "noop": {},
            "move": {
                "allowed_directions": ["north", "south", "west", "east", "northwest", "northeast", "southwest", "southeast"],
                "consumed_resources": {"energy": 1},
            },
            "change_vibe": {"vibes": _vibes(VIBES)},
            "attack": {
                "enabled": True,
                "vibes": ["weapon"],
                "consumed_resources": {"weapon": 1},
                "defense_resources": {"shield": 1},
                "armor_resources": {"shield": 1},
                "weapon_resources": {"weapon": 1},
                "vibe_bonus": {"shield": 2},
                "success": {"actor_inv_delta": {"battery": 1}, "target_inv_delta": {"energy": -7}, "loot": ["heart", "carbon"], "freeze": 3},
            },
            "transfer": {
                "enabled": True,
                "vibe_transfers": [
                    {"vibe": "charger", "target": {"energy": 10}, "actor": {"energy": -10}},
                    {"vibe": "heart", "target": {"heart": 1, "carbon": -1}, "actor": {"heart": -1, "carbon": 1}},
                ],
            },
        },
        "objects": {
            "wall": {"pydantic_type": "wall", "name": "wall"},
            "block": {"pydantic_type": "wall", "name": "block", "vibe": 2, "tags": ["block", "heavy"]},
            "extractor": {
                "pydantic_type": "assembler",
                "name": "extractor",
                "protocols": [{"output_resources": {"carbon": 3, "oxygen": 2}, "cooldown": 6}],
                "allow_partial_usage": True,
            },
            "forge": {
                "pydantic_type": "assembler",
                "name": "forge",
                "max_uses": 5,
                "chest_search_distance": 2,
                "protocols": [
                    {"input_resources": {"carbon": 2}, "output_resources": {"battery": 1}, "cooldown": 2},
                    {"vibes": ["heart"], "input_resources": {"carbon": 1, "oxygen": 1}, "output_resources": {"heart": 2}, "cooldown": 3},
                    {"vibes": ["heart", "gear"], "input_resources": {"carbon": 2, "energy": 5}, "output_resources": {"heart": 5, "gear": 1}, "cooldown": 4},
                    {"vibes": ["carbon"], "min_agents": 2, "input_resources": {"energy": 3}, "output_resources": {"shield": 1, "weapon": 1}, "cooldown": 1},
                ],
            },
            "chest": {
                "pydantic_type": "chest",
                "name": "chest",
                "vibe_transfers": {
                    "default": {"heart": 2, "carbon": -1},
                    "carbon": {"carbon": 5},
                    "heart": {"heart": -3, "oxygen": 1},
                },
                "inventory": {"initial": {"carbon": 6, "heart": 1}, "limits": {"h": {"limit": 9, "resources": ["heart"]}}},
            },
        },
    }
    legend = {"#": "wall", ".": "empty", "B": "block", "E": "extractor", "F": "forge", "C": "chest", "a": "agent.team_0", "b": "agent.team_1"}
    rows = [
        "############",
        "#a.E..b...B#",
        "#..a.F.b..##",
        "#.C..ab.E..#",
        "#....F.C...#",
        "#..B......a#",
        "#b..E...F..#",
        "############",
    ]
    return MettaGridConfig(game=game), ascii_map(rows, legend)


def tiny_truncating() -> tuple[MettaGridConfig, np.ndarray]:
    """2 agents, max_steps=5, truncation: drives the auto-reset path (MGT/test_mettagrid_puffer_env.py:208-251)."""
    game = {
        "resource_names": ["heart", "ore"],
        "num_agents": 2,
        "max_steps": 5,
        "episode_truncates": True,
        "obs": {"width": 5, "height": 5, "num_tokens": 30},
        "agent": {"rewards": {"inventory": {"heart": 1.0}}, "inventory": {"default_limit": 10}},
        "actions": {"noop": {}, "move": {}, "change_vibe": {"enabled": False}},
        "objects": {
            "wall": {"pydantic_type": "wall", "name": "wall"},
            "altar": {"pydantic_type": "assembler", "name": "altar", "protocols": [{"output_resources": {"heart": 1}, "cooldown": 2}]},
        },
    }
    legend = {"#": "wall", ".": "empty", "A": "altar", "@": "agent.agent"}
    rows = ["#######", "#@.A..#", "#..@..#", "#######"]
    return MettaGridConfig(game=game), ascii_map(rows, legend)


def clipped_start() -> tuple[MettaGridConfig, np.ndarray]:
    """start_clipped assemblers with a clipper whose clip_period is 0: deterministic in the reference."""
    game = {
        "resource_names": ["energy", "carbon", "decoder", "modulator", "heart"],
        "num_agents": 3,
        "max_steps": 0,
        "obs": {"width": 11, "height": 11, "num_tokens": 80},
        "agent": {"inventory": {"default_limit": 50, "initial": {"decoder": 2, "modulator": 2, "energy": 20}}, "rewards": {"inventory": {"heart": 1.0}}},
        "actions": {"noop": {}, "move": {}, "change_vibe": {"vibes": _vibes(["default", "gear"])}},
        "clipper": {
            "unclipping_protocols": [
                {"input_resources": {"decoder": 1}, "cooldown": 1},
                {"input_resources": {"modulator": 1}, "cooldown": 1},
            ],
            "clip_period": 0,
        },
        "objects": {
            "wall": {"pydantic_type": "wall", "name": "wall"},
            "mine": {"pydantic_type": "assembler", "name": "mine", "start_clipped": True,
                     "protocols": [{"output_resources": {"carbon": 1}, "cooldown": 3}]},
            "altar": {"pydantic_type": "assembler", "name": "altar", "clip_immune": True,
                      "protocols": [{"input_resources": {"carbon": 1}, "output_resources": {"heart": 1}, "cooldown": 1}]},
        },
    }
    legend = {"#": "wall", ".": "empty", "M": "mine", "A": "altar", "@": "agent.agent"}
    rows = ["#########", "#@M...M.#", "#..@A...#", "#.M...@.#", "#########"]
    return MettaGridConfig(game=game), ascii_map(rows, legend)


def _crowded(num_agents: int, rows: list[str]) -> tuple[MettaGridConfig, np.ndarray]:
    """kitchen_sink's rules in a room where nearly every action meets another agent: the processing
    order of the reference (std::shuffle per step) decides most outcomes. Loot can take the `gear`
    item that raises the cargo limit, moves consume nothing (agents never run dry and stop), attacks
    freeze for a few steps so that swaps with frozen agents happen all the time."""
    cfg, _ = kitchen_sink()
    game = cfg.game.model_dump()
    game["num_agents"] = num_agents
    teams = [i % 2 for i in range(num_agents)]
    proto = game["agents"][0]
    game["agents"] = []
    for t in teams:
        a = {k: v for k, v in proto.items()}
        a["team_id"] = t
        a["tags"] = ["agent", f"team{t}"]
        a["freeze_duration"] = 3
        game["agents"].append(a)
    game["actions"]["move"]["consumed_resources"] = {}
    game["actions"]["attack"]["consumed_resources"] = {}
    game["actions"]["attack"]["success"]["loot"] = ["heart", "carbon", "gear"]
    game["actions"]["attack"]["success"]["freeze"] = 2
    game["inventory_regen_interval"] = 2
    legend = {"#": "wall", ".": "empty", "B": "block", "E": "extractor", "F": "forge", "C": "chest", "a": "agent.team_0", "b": "agent.team_1"}
    return MettaGridConfig(game=game), ascii_map(rows, legend)


def _populate(rows: list[str], num_agents: int, seed: int) -> list[str]:
    """Scatter `num_agents` agents (teams alternating) over the free cells of an object-only room, reproducibly."""
    cells = [(r, c) for r, row in enumerate(rows) for c, ch in enumerate(row) if ch == "."]
    pick = np.random.RandomState(seed).permutation(len(cells))[:num_agents]
    out = [list(row) for row in rows]
    for k, idx in enumerate(sorted(pick)):
        r, c = cells[idx]
        out[r][c] = "ab"[k % 2]
    return ["".join(row) for row in out]


def dense_conflict() -> tuple[MettaGridConfig, np.ndarray]:
    """28 agents in an 8x7 room with two forges (chest search distance 2, a multi-vibe protocol with
    min_agents 2), an extractor and two chests."""
    rows = [
        "##########",
        "#........#",
        "#..F...C.#",
        "#........#",
        "#...E..F.#",
        "#.C......#",
        "#....B...#",
        "#........#",
        "##########",
    ]
    return _crowded(28, _populate(rows, 28, seed=1))


def crowd_48() -> tuple[MettaGridConfig, np.ndarray]:
    """48 agents (more than a warp has lanes: the step kernel's 64-agent instantiation) in a 12x12 room."""
    rows = [
        "##############",
        "#............#",
        "#..F......C..#",
        "#............#",
        "#....E.......#",
        "#............#",
        "#......F.....#",
        "#.C..........#",
        "#............#",
        "#.....E....B.#",
        "#............#",
        "#..C.....F...#",
        "#............#",
        "##############",
    ]
    return _crowded(48, _populate(rows, 48, seed=2))


def clipper_spread() -> tuple[MettaGridConfig, np.ndarray]:
    """A clipper that actually picks: 14 trackable assemblers spread over a 17x21 map, clip_period 6, one
    assembler starting clipped (so infection weights exist from step 0), cheap unclipping protocols and agents
    holding the items for them (random walkers unclip now and then, which moves assemblers between the
    clipper's border / unclipped sets): Clipper::pick_assembler_to_clip, clip_assembler, on_unclip_assembler
    (systems/clipper.hpp:140-238)."""
    game = {
        "resource_names": ["energy", "carbon", "decoder", "modulator", "heart"],
        "num_agents": 6,
        "max_steps": 0,
        "obs": {"width": 11, "height": 11, "num_tokens": 100},
        "agent": {"inventory": {"default_limit": 200, "initial": {"decoder": 60, "modulator": 60, "energy": 20}}, "rewards": {"inventory": {"heart": 1.0}}},
        "actions": {"noop": {}, "move": {}, "change_vibe": {"vibes": _vibes(["default", "gear"])}},
        "clipper": {
            "unclipping_protocols": [
                {"input_resources": {"decoder": 1}, "cooldown": 1},
                {"input_resources": {"modulator": 1}, "cooldown": 2},
                {"input_resources": {"energy": 1}, "cooldown": 1},
            ],
            "clip_period": 6,
            "scaled_cutoff_distance": 2,
        },
        "objects": {
            "wall": {"pydantic_type": "wall", "name": "wall"},
            "mine": {"pydantic_type": "assembler", "name": "mine", "protocols": [{"output_resources": {"carbon": 1}, "cooldown": 3}]},
            "seed": {"pydantic_type": "assembler", "name": "seed", "start_clipped": True,
                     "protocols": [{"output_resources": {"energy": 1}, "cooldown": 2}]},
            "altar": {"pydantic_type": "assembler", "name": "altar", "clip_immune": True,
                      "protocols": [{"input_resources": {"carbon": 1}, "output_resources": {"heart": 1}, "cooldown": 1}]},
        },
    }
    legend = {"#": "wall", ".": "empty", "M": "mine", "S": "seed", "A": "altar", "@": "agent.agent"}
    rows = [
        "#####################",
        "#.M....@...M......M.#",
        "#...................#",
        "#....M.....@....M...#",
        "#.@.......S.........#",
        "#...M..........M....#",
        "#.........A.....@...#",
        "#..M......M.........#",
        "#.....@........M....#",
        "#...........M.......#",
        "#.M.......@......M..#",
        "#####################",
    ]
    return MettaGridConfig(game=game), ascii_map(rows, legend)


def collective_rewards() -> tuple[MettaGridConfig, np.ndarray]:
    """Collectives as the core uses them without activation handlers (objects/collective.hpp, agent.cpp:115-117):
    membership by "collective:<name>" tag, reward terms that read the collective's stats (aligned.agent,
    collective.<res>.deposited of its initial inventory), plus reward terms on objects.<cell name> game stats."""
    def agent(team, coll):
        return {
            "team_id": team,
            "tags": ["agent", f"collective:{coll}"],
            "inventory": {"default_limit": 20, "initial": {"heart": 1}},
            "rewards": {
                "inventory": {"heart": 0.5},
                "stats": {"aligned.agent": 0.25, "collective.heart.deposited": 0.125, "collective.ore.deposited": 2.0,
                          "objects.wall": 0.0078125, "objects.altar": 0.5, "action.invalid_index": 0.0625, "objects.nothing": 3.0},
                "stats_max": {"objects.wall": 0.25},
            },
        }

    game = {
        "resource_names": ["heart", "ore"],
        "num_agents": 5,
        "max_steps": 0,
        "obs": {"width": 5, "height": 5, "num_tokens": 40},
        "agents": [agent(0, "red"), agent(0, "red"), agent(0, "red"), agent(1, "blue"), agent(1, "blue")],
        "collectives": [{"name": "red", "inventory": {"initial": {"heart": 6}}}, {"name": "blue", "inventory": {"initial": {"ore": 3, "heart": 0}}}],
        "actions": {"noop": {}, "move": {}, "change_vibe": {"enabled": False}},
        "objects": {
            "wall": {"pydantic_type": "wall", "name": "wall"},
            "altar": {"pydantic_type": "assembler", "name": "altar", "max_uses": 7, "protocols": [{"output_resources": {"heart": 1}, "cooldown": 2}]},
        },
    }
    legend = {"#": "wall", ".": "empty", "A": "altar", "a": "agent.team_0", "b": "agent.team_1"}
    rows = ["########", "#a.A..b#", "#..a.A.#", "#b...a.#", "########"]
    return MettaGridConfig(game=game), ascii_map(rows, legend)


SCENARIOS = {"kitchen_sink": kitchen_sink, "tiny_truncating": tiny_truncating, "clipped_start": clipped_start,
             "dense_conflict": dense_conflict, "crowd_48": crowd_48, "clipper_spread": clipper_spread, "collective_rewards": collective_rewards}
