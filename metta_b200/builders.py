"""Game content: the reference's stock environment configs, expressed in our schema mirror.

Reference: packages/mettagrid/python/src/mettagrid/builder/envs.py:26-130 (make_arena,
make_navigation) and builder/building.py:7-64 (assembler recipes). Only the game rules are
restated; map building is in metta_b200.maps.
"""

from __future__ import annotations

from .config import MettaGridConfig


def _assembler(name: str, inputs: dict[str, int], outputs: dict[str, int], cooldown: int) -> dict:
    return {
        "pydantic_type": "assembler",
        "name": name,
        "protocols": [{"input_resources": inputs, "output_resources": outputs, "cooldown": cooldown}],
    }


def arena_objects() -> dict:
    # building.py:10-56
    return {
        "wall": {"pydantic_type": "wall", "name": "wall"},
        "assembler": _assembler("assembler", {"battery_red": 3}, {"heart": 1}, 10),
        "mine_red": _assembler("mine_red", {}, {"ore_red": 1}, 50),
        "generator_red": _assembler("generator_red", {"ore_red": 1}, {"battery_red": 1}, 25),
        "lasery": _assembler("lasery", {"battery_red": 1, "ore_red": 2}, {"laser": 1}, 10),
        "armory": _assembler("armory", {"ore_red": 3}, {"armor": 1}, 10),
    }


def make_arena(num_agents: int = 24, combat: bool = True, max_steps: int = 10000) -> MettaGridConfig:
    """builder/envs.py:26-103. 5 actions (noop + 4 moves); attack is configured but no vibe triggers it."""
    game = {
        "num_agents": num_agents,
        "max_steps": max_steps,
        "actions": {
            "noop": {},
            "move": {},
            "attack": {"consumed_resources": {"laser": 1 if combat else 100}, "defense_resources": {"armor": 1}},
            "change_vibe": {"enabled": False},
        },
        "objects": arena_objects(),
        "agent": {
            "inventory": {"default_limit": 50, "limits": {"heart": {"limit": 255, "resources": ["heart"]}}},
            "rewards": {"inventory": {"heart": 1}},
        },
    }
    return MettaGridConfig(label="arena" + (".combat" if combat else ""), game=game)


def make_navigation(num_agents: int = 4, max_steps: int = 10000) -> MettaGridConfig:
    """builder/envs.py:106-130: walls plus heart-dispensing assemblers, one resource."""
    game = {
        "num_agents": num_agents,
        "max_steps": max_steps,
        "resource_names": ["heart"],
        "objects": {
            "assembler": _assembler("assembler", {}, {"heart": 1}, 255),
            "wall": {"pydantic_type": "wall", "name": "wall"},
        },
        "actions": {"move": {"enabled": True}, "noop": {"enabled": True}},
        "agent": {"rewards": {"inventory": {"heart": 1}}},
    }
    return MettaGridConfig(game=game)
