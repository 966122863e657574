"""GameConfig -> lowered int32 blob (layout: include/mg_config.h).

This is the counterpart of the reference's ``convert_to_cpp_game_config``
(packages/mettagrid/python/src/mettagrid/config/mettagrid_c_config.py:30-577): the same id
assignment rules (resource ids, type ids, vibe ids, tag ids, feature ids, action order), but the
result is a flat array for the kernels instead of a tree of pybind objects. Every list whose
order the reference inherits from a ``std::unordered_map`` is ordered here by asking the real
container (csrc/mg_host_order.cpp), so device code never hashes anything.
"""

from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import Iterable

import numpy as np

from .config import AssemblerConfig, ChestConfig, GameConfig, InventoryConfig, ProtocolConfig, WallConfig
from .layout import L, list_item, list_val

ORIENTATIONS = {
    # orientation.hpp:6-15
    "north": 0,
    "south": 1,
    "west": 2,
    "east": 3,
    "northwest": 4,
    "northeast": 5,
    "southwest": 6,
    "southeast": 7,
}
TEAM_NAMES = {0: "red", 1: "blue", 2: "green", 3: "yellow", 4: "purple", 5: "orange"}  # mettagrid_c_config.py:148


def f32_bits(x: float) -> int:
    return struct.unpack("<i", struct.pack("<f", float(x)))[0]


# ---------------------------------------------------------------- unordered_map order helpers


def _umap_order_u8(keys: list[int], reserve_first: bool) -> list[int]:
    """Positions of ``keys`` in the iteration order of unordered_map<uint8_t,..> after inserting
    them in the given order (``reserve_first``: pybind11's loader reserves len(dict) first)."""
    if len(keys) <= 1:
        return list(range(len(keys)))
    from .capi import lib

    k = np.asarray(keys, dtype=np.uint8)
    out = np.zeros(len(keys), dtype=np.int32)
    n = lib().mg_umap_order_u8(k.ctypes.data, len(keys), int(reserve_first), out.ctypes.data)
    assert n == len(keys), "duplicate keys in a config map"
    return [int(i) for i in out]


def _umap_order_str(keys: list[str], reserve_first: bool) -> list[int]:
    if len(keys) <= 1:
        return list(range(len(keys)))
    from .capi import lib

    buf = b"".join(s.encode() + b"\0" for s in keys)
    out = np.zeros(len(keys), dtype=np.int32)
    n = lib().mg_umap_order_str(buf, len(keys), int(reserve_first), out.ctypes.data)
    assert n == len(keys)
    return [int(i) for i in out]


def map_order(pairs: list[tuple[int, int]], reserve_first: bool = True) -> list[tuple[int, int]]:
    """(item, value) pairs of a Python dict -> the order the C++ map iterates them in."""
    order = _umap_order_u8([p[0] for p in pairs], reserve_first)
    return [pairs[i] for i in order]


# ---------------------------------------------------------------- result container


@dataclass
class LoweredConfig:
    words: np.ndarray  # int32 blob
    cell_codes: dict[str, int]  # map cell name -> code (1-based; empty names -> 0)
    action_names: list[str]
    resource_names: list[str]
    vibe_names: list[str]
    tag_names: list[str]
    feature_ids: dict[str, int]
    feature_normalizations: dict[int, float]
    type_names: dict[int, str]  # reference type_id -> type name
    group_names: list[str]  # per agent config
    cell_kind: dict[int, int] = field(default_factory=dict)  # code -> MG_KIND_*
    cell_names: dict[int, str] = field(default_factory=dict)  # code -> first map name
    agent_stat_names: dict[str, int] = field(default_factory=dict)  # readable name -> slot (for group 0..)
    num_agents: int = 0

    def __getitem__(self, off: int) -> int:
        return int(self.words[off])


class _Blob:
    def __init__(self) -> None:
        self.w: list[int] = [0] * L.MGC_HEADER_WORDS

    def alloc(self, n: int) -> int:
        off = len(self.w)
        self.w.extend([0] * n)
        return off

    def put_list(self, off: int, pairs: Iterable[tuple[int, int]]) -> None:
        pairs = list(pairs)
        if len(pairs) > L.MG_RES_STRIDE:
            raise ValueError("item list longer than MG_RES_STRIDE")
        self.w[off + L.MG_LIST_N] = len(pairs)
        for i, (item, val) in enumerate(pairs):
            self.w[off + list_item(i)] = int(item)
            self.w[off + list_val(i)] = int(val)


def num_inventory_tokens(base: int) -> int:
    """observation_encoder.hpp:83-97 compute_num_tokens(65535, base)."""
    if base <= 1:
        raise ValueError("Base must be greater than 1")
    n, v = 0, 65535
    while v > 0:
        v //= base
        n += 1
    return n


def compute_features(game: GameConfig) -> list[tuple[int, str, float]]:
    """Feature id table; same order as IdMap._compute_features (config/id_map.py:91-180)."""
    feats: list[tuple[str, float]] = [
        ("agent:group", 10.0),
        ("agent:frozen", 1.0),
        ("episode_completion_pct", 255.0),
        ("last_action", 10.0),
        ("last_reward", 100.0),
        ("goal", 100.0),
        ("vibe", 255.0),
        ("agent:compass", 1.0),
        ("tag", 10.0),
        ("cooldown_remaining", 255.0),
        ("clipped", 1.0),
        ("remaining_uses", 255.0),
    ]
    base = game.obs.token_value_base
    # id_map.py:23-36 uses ceil(log(65536, base)); identical to the integer loop for base >= 2
    ntok = num_inventory_tokens(base)
    for r in game.resource_names:
        feats.append((f"inv:{r}", float(base)))
        for p in range(1, ntok):
            feats.append((f"inv:{r}:p{p}", float(base)))
    if game.protocol_details_obs:
        for r in game.resource_names:
            feats.append((f"protocol_input:{r}", 100.0))
        for r in game.resource_names:
            feats.append((f"protocol_output:{r}", 100.0))
    return [(i, n, norm) for i, (n, norm) in enumerate(feats)]


def _invcfg(blob: _Blob, base: int, rid: dict[str, int], inv: InventoryConfig, with_defaults: bool, R: int) -> None:
    """Write one INVCFG record. Agents get default limits for unconfigured resources
    (mettagrid_c_config.py:203-224); chests only explicit limits (:323-339)."""
    w = blob.w
    limit_defs: list[tuple[list[int], int, dict[int, int]]] = []
    configured: set[str] = set()
    for rl in inv.limits.values():
        if with_defaults:
            ids = [rid[n] for n in rl.resources]
        else:
            ids = [rid[n] for n in rl.resources if n in rid]
            if not ids:
                continue
        mods = {rid[n]: b for n, b in rl.modifiers.items() if n in rid}
        limit_defs.append((ids, rl.limit, mods))
        configured.update(rl.resources)
    if with_defaults:
        for n in rid:  # resource_names order
            if n not in configured:
                limit_defs.append(([rid[n]], inv.default_limit, {}))
    if len(limit_defs) > L.MG_MAX_LIMITS:
        raise ValueError(f"more than {L.MG_MAX_LIMITS} inventory limit groups")
    for i in range(L.MG_RES_STRIDE):
        w[base + L.MG_INVCFG_ITEM_LIMIT + i] = -1
    insert_order: list[int] = []
    for li, (ids, limit, mods) in enumerate(limit_defs):  # inventory.cpp:14-24: later defs win
        w[base + L.MG_INVCFG_LIMIT_BASE + li] = int(limit) & 0xFFFF
        for item, bonus in mods.items():
            w[base + L.MG_INVCFG_LIMIT_MOD + li * L.MG_RES_STRIDE + item] = int(bonus) & 0xFFFF
        for item in ids:
            w[base + L.MG_INVCFG_ITEM_LIMIT + item] = li
            if item not in insert_order:
                insert_order.append(item)
    w[base + L.MG_INVCFG_NUM_LIMITS] = len(limit_defs)
    reachable = {w[base + L.MG_INVCFG_ITEM_LIMIT + i] for i in range(R)} - {-1}
    for li in reachable:  # inventory.cpp:119-126 is_modifier: key present, bonus may be 0
        for item in limit_defs[li][2]:
            w[base + L.MG_INVCFG_IS_MODIFIER + item] = 1
    order = _umap_order_u8(insert_order, reserve_first=False)  # _limits is default-constructed
    w[base + L.MG_INVCFG_NUM_ORDER] = len(order)
    for i, pos in enumerate(order):
        w[base + L.MG_INVCFG_ORDER + i] = insert_order[pos]


def _stat_slots(name: str, rid: dict[str, int], own_group: str, group_index: dict[str, int],
                cell_codes: dict[str, int] | None = None) -> tuple[int, int]:
    """Resolve a stat name used as a reward key to (agent slot, game slot), -1 = never emitted.

    Names: SURVEY.md Appendix C (action_handler.hpp:102-149, agent.cpp:78-99,171-175,
    chest.hpp:41-62, assembler.hpp:181, mettagrid_c.cpp:552-569, attack.hpp:263-275,
    transfer.hpp:148,174).
    """
    parts = name.split(".")
    a = g = -1
    simple = {
        "action.noop.success": L.MG_AS_ACTION_SUCCESS + 0,
        "action.move.success": L.MG_AS_ACTION_SUCCESS + 1,
        "action.change_vibe.success": L.MG_AS_ACTION_SUCCESS + 2,
        "action.noop.failed": L.MG_AS_ACTION_FAILED + 0,
        "action.move.failed": L.MG_AS_ACTION_FAILED + 1,
        "action.change_vibe.failed": L.MG_AS_ACTION_FAILED + 2,
        "action.failed": L.MG_AS_ACTION_FAILED_ANY,
        "status.frozen.ticks": L.MG_AS_FROZEN_TICKS,
        f"status.frozen.ticks.{own_group}": L.MG_AS_FROZEN_TICKS,
        "status.max_steps_without_motion": L.MG_AS_MAX_STEPS_WITHOUT_MOTION,
        "actions.swap": L.MG_AS_SWAP,
        "inventory.diversity": L.MG_AS_DIVERSITY,
        "damage.items_lost": L.MG_AS_DAMAGE_ITEMS_LOST,
        "damage.triggered": L.MG_AS_DAMAGE_TRIGGERED,
        "action.invalid_index": L.MG_AS_INVALID_INDEX,
        f"action.attack.{own_group}.friendly_fire": L.MG_AS_ATTACK_FRIENDLY_FIRE,
        f"action.transfer.{own_group}.count": L.MG_AS_TRANSFER_COUNT,
    }
    game_simple = {
        "tokens_written": L.MG_GS_TOKENS_WRITTEN,
        "tokens_dropped": L.MG_GS_TOKENS_DROPPED,
        "tokens_free_space": L.MG_GS_TOKENS_FREE_SPACE,
    }
    if name in simple:
        a = simple[name]
    if name in game_simple:
        g = game_simple[name]
    if len(parts) == 2 and parts[0] in rid:
        r = rid[parts[0]]
        a = {"amount": L.MG_AS_AMOUNT + r, "gained": L.MG_AS_GAINED + r, "lost": L.MG_AS_LOST + r}.get(parts[1], -1)
    if len(parts) == 2 and parts[0] == "damaged" and parts[1] in rid:
        a = L.MG_AS_DAMAGED + rid[parts[1]]
    if len(parts) == 3 and parts[0] == "chest" and parts[1] in rid:
        r = rid[parts[1]]
        a = {"deposited_by_agent": L.MG_AS_CHEST_DEPOSITED + r, "withdrawn_by_agent": L.MG_AS_CHEST_WITHDRAWN + r}.get(
            parts[2], -1
        )
        g = {
            "deposited": L.MG_GS_CHEST_DEPOSITED + r,
            "withdrawn": L.MG_GS_CHEST_WITHDRAWN + r,
            "amount": L.MG_GS_CHEST_AMOUNT + r,
        }.get(parts[2], -1)
    if len(parts) == 3 and parts[0] == "assembler" and parts[1] in rid and parts[2] == "created":
        g = L.MG_GS_ASM_CREATED + rid[parts[1]]
    if name.startswith(f"action.attack.{own_group}."):
        rest = name[len(f"action.attack.{own_group}.") :]
        for prefix, base in (("blocked_by.", L.MG_AS_ATTACK_BLOCKED_BY), ("hit.", L.MG_AS_ATTACK_HIT)):
            if rest.startswith(prefix) and rest[len(prefix) :] in group_index:
                a = base + group_index[rest[len(prefix) :]]
        if rest.startswith("hit_by.") and rest[len("hit_by.") :] in group_index:
            a = L.MG_AS_ATTACK_HIT_BY + group_index[rest[len("hit_by.") :]]
    if name.startswith(f"action.transfer.{own_group}."):
        rest = name[len(f"action.transfer.{own_group}.") :].split(".")
        if len(rest) == 2 and rest[0] in rid and rest[1] == "self":
            a = L.MG_AS_TRANSFER_SELF + rid[rest[0]]
        if len(rest) == 3 and rest[0] in rid and rest[1] == "to" and rest[2] in group_index:
            a = L.MG_AS_TRANSFER_TO + group_index[rest[2]] * L.MG_RES_STRIDE + rid[rest[0]]
    if name.startswith("objects.") and cell_codes and name[len("objects.") :] in cell_codes:
        g = L.MG_GS_OBJECTS + cell_codes[name[len("objects.") :]]  # counted per map when an episode starts
    return a, g


def _lower_protocol(blob: _Blob, off: int, p: ProtocolConfig, rid: dict[str, int], vid: dict[str, int], where: str):
    w = blob.w
    for v in p.vibes:
        if v not in vid:
            raise ValueError(f"Unknown vibe name '{v}' in {where}")
    vibes = sorted(vid[v] for v in p.vibes)
    if len(vibes) > L.MG_MAX_PROTO_VIBES:
        raise ValueError("protocol with more than 8 vibes")
    key = 0
    for v in vibes:  # assembler.hpp:281-288
        key = ((key << 8) | v) & 0xFFFFFFFFFFFFFFFF
    w[off + L.MG_PROTO_KEY_LO] = _i32(key & 0xFFFFFFFF)
    w[off + L.MG_PROTO_KEY_HI] = _i32(key >> 32)
    w[off + L.MG_PROTO_MIN_AGENTS] = int(p.min_agents)
    w[off + L.MG_PROTO_NUM_VIBES] = len(vibes)
    for i, v in enumerate(vibes):
        w[off + L.MG_PROTO_VIBES + i] = v
    w[off + L.MG_PROTO_COOLDOWN] = int(p.cooldown) & 0xFFFF
    inputs = map_order([(rid[k], int(v)) for k, v in p.input_resources.items()])
    outputs = map_order([(rid[k], int(v)) for k, v in p.output_resources.items()])
    blob.put_list(off + L.MG_PROTO_INPUTS, inputs)
    blob.put_list(off + L.MG_PROTO_OUTPUTS, outputs)
    # scale_protocol_for_partial_usage (assembler.hpp:427-453) rebuilds both maps with operator[]
    # in iteration order of the original: a default-constructed map, no reserve.
    blob.put_list(off + L.MG_PROTO_SCALED_INPUTS, map_order(inputs, reserve_first=False))
    blob.put_list(off + L.MG_PROTO_SCALED_OUTPUTS, map_order(outputs, reserve_first=False))
    w[off + L.MG_PROTO_HAS_OUTPUT] = int(any(v > 0 for _, v in outputs))
    return key, int(p.min_agents)


def _i32(x: int) -> int:
    x &= 0xFFFFFFFF
    return x - (1 << 32) if x & 0x80000000 else x


def lower_game_config(
    game: GameConfig, map_h: int, map_w: int, max_assemblers: int = 0, max_chests: int = 0
) -> LoweredConfig:
    resource_names = list(game.resource_names)
    R = len(resource_names)
    if R > L.MG_MAX_RESOURCES:
        raise ValueError(f"at most {L.MG_MAX_RESOURCES} resources are supported, got {R}")
    rid = {n: i for i, n in enumerate(resource_names)}
    vibes = game.actions.change_vibe.vibes
    vibe_names = [v.name for v in vibes]
    if len(vibe_names) > L.MG_MAX_VIBES:
        raise ValueError("too many vibes")
    vid = {n: i for i, n in enumerate(vibe_names)}
    type_id = {n: i + 1 for i, n in enumerate(sorted(game.objects.keys()))}  # mettagrid_c_config.py:54-55

    agents = list(game.agents)
    if not agents:  # mettagrid_c_config.py:69-77
        agents = [game.agent.model_copy(update={"team_id": 0}) for _ in range(game.num_agents)]
    all_tags: set[str] = set()
    for oc in game.objects.values():
        all_tags.update(oc.tags)
    for ac in agents:
        all_tags.update(ac.tags)
    tag_names = sorted(all_tags)
    if len(tag_names) > 256:
        raise ValueError(f"Too many unique tags ({len(tag_names)}). Maximum supported is 256 due to uint8 limit.")
    tid = {t: i for i, t in enumerate(tag_names)}

    feats = compute_features(game)
    fid = {n: i for i, n, _ in feats}
    if len(feats) > 255:
        raise ValueError("more than 255 observation features")
    ntok = num_inventory_tokens(game.obs.token_value_base)
    if ntok > L.MG_MAX_INV_TOKENS:
        raise ValueError("token_value_base too small")
    if game.obs.width > 15 or game.obs.height > 15:  # mettagrid_c.cpp:65-70
        raise ValueError(
            f"Observation window size ({game.obs.width}x{game.obs.height}) exceeds maximum packable size"
        )

    blob = _Blob()
    w = blob.w
    w[L.MGC_MAGIC] = L.MG_CONFIG_MAGIC
    w[L.MGC_VERSION] = L.MG_CONFIG_VERSION
    w[L.MGC_NUM_AGENTS] = game.num_agents
    w[L.MGC_MAP_H], w[L.MGC_MAP_W] = map_h, map_w
    w[L.MGC_OBS_H], w[L.MGC_OBS_W] = game.obs.height, game.obs.width
    w[L.MGC_NUM_TOKENS] = game.obs.num_tokens
    w[L.MGC_TOKEN_BASE] = game.obs.token_value_base
    w[L.MGC_NUM_INV_TOKENS] = ntok
    w[L.MGC_MAX_STEPS] = game.max_steps
    w[L.MGC_EPISODE_TRUNCATES] = int(game.episode_truncates)
    w[L.MGC_REGEN_INTERVAL] = game.inventory_regen_interval
    w[L.MGC_NUM_RESOURCES] = R
    w[L.MGC_PROTOCOL_DETAILS_OBS] = int(game.protocol_details_obs)
    g = game.global_obs
    w[L.MGC_GOBS_ECP] = int(g.episode_completion_pct)
    w[L.MGC_GOBS_LAST_ACTION] = int(g.last_action)
    w[L.MGC_GOBS_LAST_REWARD] = int(g.last_reward)
    w[L.MGC_GOBS_COMPASS] = int(g.compass)
    w[L.MGC_GOBS_GOAL] = int(g.goal_obs)
    for key, name in (
        (L.MGC_FEAT_GROUP, "agent:group"),
        (L.MGC_FEAT_FROZEN, "agent:frozen"),
        (L.MGC_FEAT_ECP, "episode_completion_pct"),
        (L.MGC_FEAT_LAST_ACTION, "last_action"),
        (L.MGC_FEAT_LAST_REWARD, "last_reward"),
        (L.MGC_FEAT_VIBE, "vibe"),
        (L.MGC_FEAT_COMPASS, "agent:compass"),
        (L.MGC_FEAT_TAG, "tag"),
        (L.MGC_FEAT_COOLDOWN, "cooldown_remaining"),
        (L.MGC_FEAT_CLIPPED, "clipped"),
        (L.MGC_FEAT_REMAINING_USES, "remaining_uses"),
        (L.MGC_FEAT_GOAL, "goal"),
    ):
        w[key] = fid[name]
    w[L.MGC_NUM_AGENT_STATS] = L.MG_AS_COUNT
    w[L.MGC_NUM_GAME_STATS] = L.MG_GS_COUNT
    w[L.MGC_MAX_ASSEMBLERS] = max_assemblers
    w[L.MGC_MAX_CHESTS] = max_chests

    # ---- actions: the core always creates noop, then one move per allowed direction, then one
    # change_vibe per vibe (mettagrid_c.cpp:291-352); `enabled` only zeroes the vibe count.
    acts = game.actions
    action_recs: list[tuple[int, int]] = [(L.MG_ACTION_NOOP, 0)]
    action_names = ["noop"]
    for d in acts.move.allowed_directions:
        action_recs.append((L.MG_ACTION_MOVE, ORIENTATIONS[d]))
        action_names.append(f"move_{d}")
    n_vibe_actions = len(vibes) if acts.change_vibe.enabled else 0
    for i in range(n_vibe_actions):
        action_recs.append((L.MG_ACTION_CHANGE_VIBE, i))
        action_names.append(f"change_vibe_{vibe_names[i]}")
    if len(action_recs) > L.MG_MAX_ACTIONS:
        raise ValueError("too many actions")
    w[L.MGC_NUM_ACTIONS] = len(action_recs)
    off = blob.alloc(len(action_recs) * L.MG_ACT_WORDS)
    w[L.MGC_OFF_ACTIONS] = off
    for i, (k, a) in enumerate(action_recs):
        w[off + i * L.MG_ACT_WORDS + L.MG_ACT_KIND] = k
        w[off + i * L.MG_ACT_WORDS + L.MG_ACT_ARG] = a

    def action_resources(name: str, ac) -> tuple[dict[int, int], dict[int, int]]:
        # process_action_config, mettagrid_c_config.py:395-431
        if not ac.enabled:
            return {}, {}
        missing = [r for r in ac.consumed_resources if r not in rid]
        if missing:
            raise ValueError(
                f"Action '{name}' has consumed_resources {missing} that are not in resource_names. "
                "These resources will be ignored, making the action free! "
                "Either add these resources to resource_names or disable the action."
            )
        consumed = {rid[k]: int(v) for k, v in ac.consumed_resources.items()}
        src = ac.required_resources or ac.consumed_resources
        required = {rid[k]: int(v) for k, v in src.items()}
        for item, amt in consumed.items():  # action_handler.hpp:66-81
            if item not in required:
                raise ValueError(f"Consumed resource item {item} not found in required resources")
            if required[item] < amt:
                raise ValueError(
                    f"Required resources must be >= consumed resources. Item: {item} "
                    f"required: {required[item]} < consumed: {amt}"
                )
        return required, consumed

    off = blob.alloc(3 * L.MG_ACTRES_WORDS)
    w[L.MGC_OFF_ACTION_RES] = off
    for kind, (nm, ac) in enumerate((("noop", acts.noop), ("move", acts.move), ("change_vibe", acts.change_vibe))):
        req, con = action_resources(nm, ac)
        for item, amt in req.items():
            w[off + kind * L.MG_ACTRES_WORDS + L.MG_ACTRES_REQUIRED + item] = amt & 0xFFFF
        for item, amt in con.items():
            w[off + kind * L.MG_ACTRES_WORDS + L.MG_ACTRES_CONSUMED + item] = amt & 0xFFFF

    # ---- attack / transfer / vibe dispatch
    atk = acts.attack
    atk_req, atk_con = action_resources("attack", atk)
    off = blob.alloc(L.MG_ATK_WORDS)
    w[L.MGC_OFF_ATTACK] = off
    for item, amt in atk_con.items():
        w[off + L.MG_ATK_CONSUMED + item] = amt & 0xFFFF
    for k, v in atk.armor_resources.items():
        w[off + L.MG_ATK_ARMOR + rid[k]] = int(v) & 0xFFFF
    for k, v in atk.weapon_resources.items():
        w[off + L.MG_ATK_WEAPON + rid[k]] = int(v) & 0xFFFF
    w[off + L.MG_ATK_FREEZE] = int(atk.success.freeze)
    w[off + L.MG_ATK_HAS_DEFENSE] = int(len(atk.defense_resources) > 0)
    blob.put_list(off + L.MG_ATK_DEFENSE, map_order([(rid[k], int(v) & 0xFFFF) for k, v in atk.defense_resources.items()]))
    blob.put_list(off + L.MG_ATK_ACTOR_DELTA, map_order([(rid[k], int(v)) for k, v in atk.success.actor_inv_delta.items()]))
    blob.put_list(
        off + L.MG_ATK_TARGET_DELTA, map_order([(rid[k], int(v)) for k, v in atk.success.target_inv_delta.items()])
    )
    blob.put_list(off + L.MG_ATK_LOOT, [(rid[n], 0) for n in atk.success.loot])
    for v in atk.vibes:
        if v not in vid:
            raise ValueError(f"Unknown vibe name '{v}' in attack.vibes")
    for v in atk.vibe_bonus:
        if v not in vid:
            raise ValueError(f"Unknown vibe name '{v}' in attack.vibe_bonus")

    xf = acts.transfer
    if len(xf.vibe_transfers) > L.MG_MAX_TRANSFERS:
        raise ValueError("too many vibe transfers")
    off_x = blob.alloc(L.MG_XFER_EFFECTS + len(xf.vibe_transfers) * L.MG_XFER_EFFECT_WORDS)
    w[L.MGC_OFF_TRANSFER] = off_x
    w[L.MGC_NUM_TRANSFERS] = len(xf.vibe_transfers)
    for k, v in xf.required_resources.items():
        w[off_x + L.MG_XFER_REQUIRED + rid[k]] = int(v) & 0xFFFF
    xfer_of_vibe: dict[int, int] = {}
    seen: set[str] = set()
    for i, vt in enumerate(xf.vibe_transfers):
        if vt.vibe not in vid:
            raise ValueError(f"Unknown vibe name '{vt.vibe}' in transfer.vibe_transfers")
        if vt.vibe in seen:
            raise ValueError(f"Duplicate vibe name '{vt.vibe}' in transfer.vibe_transfers")
        seen.add(vt.vibe)
        xfer_of_vibe[vid[vt.vibe]] = i
        e = off_x + L.MG_XFER_EFFECTS + i * L.MG_XFER_EFFECT_WORDS
        blob.put_list(e + L.MG_XFER_ACTOR, map_order([(rid[k], int(v)) for k, v in vt.actor.items()]))
        blob.put_list(e + L.MG_XFER_TARGET, map_order([(rid[k], int(v)) for k, v in vt.target.items()]))

    off_v = blob.alloc(L.MG_MAX_VIBES * L.MG_VIBE_WORDS)
    w[L.MGC_OFF_VIBES] = off_v
    # move.hpp:45-66: iterate the {"attack","transfer"} handler map; a later handler overwrites.
    handler_names = ["attack", "transfer"]
    handler_vibes = {"attack": [vid[v] for v in atk.vibes], "transfer": list(xfer_of_vibe.keys())}
    handler_code = {"attack": L.MG_HANDLER_ATTACK, "transfer": L.MG_HANDLER_TRANSFER}
    for v in range(L.MG_MAX_VIBES):
        w[off_v + v * L.MG_VIBE_WORDS + L.MG_VIBE_TRANSFER] = xfer_of_vibe.get(v, -1)
        w[off_v + v * L.MG_VIBE_WORDS + L.MG_VIBE_RESOURCE] = -1
    for pos in _umap_order_str(handler_names, reserve_first=False):
        nm = handler_names[pos]
        for v in handler_vibes[nm]:
            w[off_v + v * L.MG_VIBE_WORDS + L.MG_VIBE_HANDLER] = handler_code[nm]
    for v, bonus in atk.vibe_bonus.items():
        w[off_v + vid[v] * L.MG_VIBE_WORDS + L.MG_VIBE_BONUS] = int(bonus)
    for v, name in enumerate(vibe_names):  # attack.hpp:171-178; vibe 0 never matches
        if v != 0 and name in rid:
            w[off_v + v * L.MG_VIBE_WORDS + L.MG_VIBE_RESOURCE] = rid[name]

    # ---- inventory configs are appended as needed
    invcfg_offsets: list[int] = []

    def add_invcfg(inv: InventoryConfig, with_defaults: bool) -> int:
        o = blob.alloc(L.MG_INVCFG_WORDS)
        _invcfg(blob, o, rid, inv, with_defaults, R)
        invcfg_offsets.append(o)
        return len(invcfg_offsets) - 1

    # ---- agent configs, one per team in order of first appearance (mettagrid_c_config.py:98-232)
    teams: dict[int, list] = {}
    for ac in agents:
        teams.setdefault(ac.team_id, []).append(ac)
    if len(teams) > L.MG_MAX_GROUPS:
        raise ValueError("too many agent teams")
    group_names = [TEAM_NAMES.get(t, f"team_{t}") for t in teams]
    group_index = {n: i for i, n in enumerate(group_names)}
    cell_defs: list[dict] = []  # index+1 = cell code
    cell_codes: dict[str, int] = {}

    def add_cell(names: list[str], kind: int, type_id_: int, param: int, vibe: int, tags: list[int]) -> None:
        if len(tags) > L.MG_MAX_TAGS:
            raise ValueError(f"more than {L.MG_MAX_TAGS} tags on {names[0]}")
        cell_defs.append(dict(kind=kind, type_id=type_id_, param=param, vibe=vibe, tags=tags, name=names[0]))
        for n in names:
            cell_codes[n] = len(cell_defs)  # later registrations win, like the dict in the reference

    # collectives (objects/collective.hpp): only agents are Alignable, and a collective's stats move only through
    # activation handlers (not lowered), so what an agent's collective adds to a reward term is a constant:
    # aligned.agent = members (addMember :36-43), collective.<res>.deposited = the initial inventory (:100-108).
    # An object joins the collective its LAST "collective:<name>" tag names (mettagrid_c.cpp:102-131).
    collective_names = {c.name for c in game.collectives}

    def collective_of(ac) -> str | None:
        name = None
        for t in ac.tags:
            if t.startswith("collective:") and t[len("collective:") :] in collective_names:
                name = t[len("collective:") :]
        return name

    collective_members = {c.name: sum(1 for ac in agents if collective_of(ac) == c.name) for c in game.collectives}
    collective_cfg = {c.name: c for c in game.collectives}

    def collective_stat(ac, stat: str) -> float:
        cname = collective_of(ac)
        if cname is None:
            return 0.0
        if stat == "aligned.agent":
            return float(collective_members[cname])
        parts = stat.split(".")
        if len(parts) == 3 and parts[0] == "collective" and parts[2] == "deposited":
            amount = collective_cfg[cname].inventory.initial.get(parts[1], 0) if parts[1] in rid else 0
            return float(amount) if amount > 0 else 0.0
        return 0.0

    pending_object_rewards: list[tuple[int, str]] = []
    group_recs: list[int] = []
    for gi, (team_id, members) in enumerate(teams.items()):
        first = members[0]
        first_tags = set(first.tags)
        for idx, other in enumerate(members[1:], start=1):
            if set(other.tags) != first_tags:
                raise ValueError(
                    f"All agents in team {team_id} must have identical tags. "
                    f"Agent 0 has tags {sorted(first_tags)}, but agent {idx} has tags {sorted(other.tags)}. "
                    "Tags are currently applied per-team, not per-agent."
                )
        go = blob.alloc(L.MG_GRP_WORDS)
        group_recs.append(go)
        gname = group_names[gi]
        w[go + L.MG_GRP_GROUP_ID] = team_id & 0xFF
        w[go + L.MG_GRP_FREEZE_DURATION] = first.freeze_duration & 0xFF  # unsigned char ctor arg
        w[go + L.MG_GRP_INITIAL_VIBE] = first.initial_vibe & 0xFF
        tags = [tid[t] for t in first.tags]
        w[go + L.MG_GRP_NUM_TAGS] = len(tags)
        for i, t in enumerate(tags[: L.MG_MAX_TAGS]):
            w[go + L.MG_GRP_TAGS + i] = t
        w[go + L.MG_GRP_INVCFG] = add_invcfg(first.inventory, with_defaults=True)
        blob.put_list(
            go + L.MG_GRP_INITIAL_INV,
            map_order([(rid[k], int(v) & 0xFFFF) for k, v in first.inventory.initial.items()]),
        )
        # rewards: stats first, then inventory rewards as "<res>.amount" (mettagrid_c_config.py:124-139)
        stat_rewards = dict(first.rewards.stats)
        stat_max = dict(first.rewards.stats_max)
        for k, v in first.rewards.inventory.items():
            assert k in rid, f"Inventory reward {k} not in resource_names"
            assert k + ".amount" not in stat_rewards, f"Stat reward {k}.amount already exists"
            stat_rewards[k + ".amount"] = v
        for k, v in first.rewards.inventory_max.items():
            assert k in rid, f"Inventory reward max {k} not in resource_names"
            assert k + ".amount" not in stat_max, f"Stat reward max {k}.amount already exists"
            stat_max[k + ".amount"] = v
        names = list(stat_rewards.keys())
        if len(names) > L.MG_MAX_REWARDS:
            raise ValueError("too many stat rewards")
        order = _umap_order_str(names, reserve_first=True)
        w[go + L.MG_GRP_NUM_REWARDS] = len(names)
        goals: list[int] = []
        goal_seen: set[str] = set()
        for i, pos in enumerate(order):
            nm = names[pos]
            a_slot, g_slot = _stat_slots(nm, rid, gname, group_index)
            ro = go + L.MG_GRP_REWARDS + i * L.MG_RWD_WORDS
            w[ro + L.MG_RWD_AGENT_STAT] = a_slot
            w[ro + L.MG_RWD_GAME_STAT] = g_slot
            w[ro + L.MG_RWD_WEIGHT] = f32_bits(stat_rewards[nm])
            w[ro + L.MG_RWD_HAS_MAX] = int(nm in stat_max)
            w[ro + L.MG_RWD_MAX] = f32_bits(stat_max.get(nm, 0.0))
            w[ro + L.MG_RWD_COLLECTIVE] = f32_bits(collective_stat(first, nm))
            if nm.startswith("objects."):
                pending_object_rewards.append((gi * L.MG_GRP_WORDS + (ro - go) + L.MG_RWD_GAME_STAT, nm[len("objects.") :]))
            # goal tokens, mettagrid_c.cpp:363-395
            dot = nm.find(".")
            if dot >= 0:
                res = nm[:dot]
                if res not in goal_seen and res in rid:
                    goals.append(fid[f"inv:{res}"])
                    goal_seen.add(res)
        if len(goals) > L.MG_MAX_GOALS:
            raise ValueError("too many goal tokens")
        w[go + L.MG_GRP_NUM_GOALS] = len(goals)
        for i, gval in enumerate(goals):
            w[go + L.MG_GRP_GOALS + i] = gval
        for n in first.diversity_tracked_resources:
            if n in rid:
                w[go + L.MG_GRP_DIVERSITY + rid[n]] = 1
        regen = first.inventory.regen_amounts
        w[go + L.MG_GRP_HAS_REGEN] = int(len(regen) > 0)
        if len(regen) > L.MG_MAX_REGEN:
            raise ValueError("too many regen vibe entries")
        w[go + L.MG_GRP_NUM_REGEN] = len(regen)
        for i, (vname, amounts) in enumerate(regen.items()):
            w[go + L.MG_GRP_REGEN_VIBE + i] = vid[vname]
            blob.put_list(
                go + L.MG_GRP_REGEN_LISTS + i * L.MG_LIST_WORDS,
                map_order([(rid[k], int(v)) for k, v in amounts.items()]),
            )
        dmg = first.damage
        if dmg is not None:
            for n in list(dmg.threshold) + list(dmg.resources):
                assert n in rid, f"Damage resource '{n}' not in resource_names"
            w[go + L.MG_GRP_DAMAGE_ENABLED] = int(bool(dmg.threshold) and bool(dmg.resources))
            blob.put_list(go + L.MG_GRP_DAMAGE_THRESHOLD, map_order([(rid[k], int(v)) for k, v in dmg.threshold.items()]))
            blob.put_list(go + L.MG_GRP_DAMAGE_RESOURCES, map_order([(rid[k], int(v)) for k, v in dmg.resources.items()]))
        names_for_cell = [f"agent.{gname}", f"agent.team_{team_id}"]
        if team_id == 0:
            names_for_cell += ["agent.default", "agent.agent"]
        add_cell(names_for_cell, L.MG_KIND_AGENT, 0, gi, first.initial_vibe, tags)
    w[L.MGC_NUM_GROUPS] = len(group_recs)

    # ---- objects
    asm_recs: list[dict] = []
    chest_recs: list[dict] = []
    protocols: list[ProtocolConfig] = []
    proto_words: list[int] = []  # offsets filled later

    def lower_protocol_list(plist: list[ProtocolConfig], where: str, sort: bool) -> tuple[int, int, list[tuple[int, int]]]:
        first = len(proto_words)
        keyed: list[tuple[int, int, int]] = []
        tmp: list[int] = []
        for p in plist:
            o = blob.alloc(L.MG_PROTO_WORDS)
            key, min_agents = _lower_protocol(blob, o, p, rid, vid, where)
            tmp.append(o)
            keyed.append((key, min_agents, o))
        if sort:
            # build_protocol_map (assembler.hpp:292-309): bucket by key in push order, then sort each
            # bucket by min_agents descending (insertion sort for short vectors: stable).
            by_key: dict[int, list[tuple[int, int, int]]] = {}
            for k in keyed:
                by_key.setdefault(k[0], []).append(k)
            keyed = [e for lst in by_key.values() for e in sorted(lst, key=lambda e: -e[1])]
        proto_words.extend(e[2] for e in keyed)
        return first, len(keyed), [(e[0], e[1]) for e in keyed]

    for object_type, oc in game.objects.items():
        tags = [tid[t] for t in oc.tags]
        names = [oc.map_name or object_type]
        if isinstance(oc, WallConfig):
            add_cell(names, L.MG_KIND_WALL, type_id[object_type], 0, oc.vibe, tags)
        elif isinstance(oc, AssemblerConfig):
            plist = list(reversed(oc.protocols))  # mettagrid_c_config.py:260
            seen_p: list[tuple[list[int], int]] = []
            for p in plist:
                for v in p.vibes:
                    if v not in vid:
                        raise ValueError(f"Unknown vibe name '{v}' in assembler '{object_type}' protocol")
                sig = (sorted(vid[v] for v in p.vibes), p.min_agents)
                if sig in seen_p:
                    raise ValueError(
                        f"Protocol with vibes {p.vibes} and min_agents {p.min_agents} already exists in {object_type}"
                    )
                seen_p.append(sig)
            first, n, keys = lower_protocol_list(plist, f"assembler '{object_type}' protocol", sort=True)
            idle = -1  # get_current_protocol with no agents around: key 0, min_agents <= 0
            for i, (k, m) in enumerate(keys):
                if k == 0 and m <= 0:
                    idle = first + i
                    break
            dependent = any(k != 0 or m > 0 for k, m in keys)
            asm_recs.append(
                dict(
                    first=first,
                    n=n,
                    partial=int(oc.allow_partial_usage),
                    max_uses=oc.max_uses,
                    immune=int(oc.clip_immune),
                    start_clipped=int(oc.start_clipped),
                    chest_distance=oc.chest_search_distance,
                    dependent=int(dependent),
                    idle=idle,
                )
            )
            add_cell(names, L.MG_KIND_ASSEMBLER, type_id[object_type], len(asm_recs) - 1, oc.vibe, tags)
        elif isinstance(oc, ChestConfig):
            vt: list[tuple[int, list[tuple[int, int]]]] = []
            for vname, deltas in oc.vibe_transfers.items():
                if vname not in vid:
                    raise ValueError(f"Unknown vibe name '{vname}' in chest '{object_type}' vibe_transfers")
                vt.append((vid[vname], map_order([(rid[k], int(v)) for k, v in deltas.items()])))
            if len(vt) > L.MG_MAX_CHEST_VIBES:
                raise ValueError("too many chest vibe transfers")
            initial = map_order([(rid[k], int(v)) for k, v in oc.inventory.initial.items()])
            chest_recs.append(dict(invcfg=add_invcfg(oc.inventory, with_defaults=False), initial=initial, vt=vt))
            add_cell(names, L.MG_KIND_CHEST, type_id[object_type], len(chest_recs) - 1, oc.vibe, tags)
        else:
            raise ValueError(f"Unknown object type: {object_type}")

    if len(asm_recs) > L.MG_MAX_ASM_TYPES or len(chest_recs) > L.MG_MAX_CHEST_TYPES:
        raise ValueError("too many assembler/chest types")
    if len(cell_defs) >= L.MG_MAX_CELLTYPES:  # codes 1..MG_MAX_CELLTYPES-1 (a game stat slot each)
        raise ValueError("too many cell types")

    # ---- clipper (mettagrid_c_config.py:512-546)
    off_c = blob.alloc(L.MG_CLIP_WORDS)
    w[L.MGC_OFF_CLIPPER] = off_c
    if game.clipper is not None:
        cl = game.clipper
        if not cl.unclipping_protocols:
            raise ValueError("Clipper config provided but unclipping_protocols is empty")  # mettagrid_c.cpp:145-147
        first, n, _ = lower_protocol_list(list(cl.unclipping_protocols), "clipper unclipping_protocols", sort=False)
        w[off_c + L.MG_CLIP_ENABLED] = 1
        w[off_c + L.MG_CLIP_FIRST_PROTOCOL] = first
        w[off_c + L.MG_CLIP_NUM_PROTOCOLS] = n
        w[off_c + L.MG_CLIP_LENGTH_SCALE] = cl.length_scale
        w[off_c + L.MG_CLIP_CUTOFF] = cl.scaled_cutoff_distance
        w[off_c + L.MG_CLIP_PERIOD] = _i32(cl.clip_period)

    # ---- sections with fixed-stride records, copied from the scratch allocations above
    if len(proto_words) > L.MG_MAX_PROTOCOLS:
        raise ValueError("too many protocols")
    off_p = blob.alloc(max(1, len(proto_words)) * L.MG_PROTO_WORDS)
    w[L.MGC_OFF_PROTOCOLS] = off_p
    w[L.MGC_NUM_PROTOCOLS] = len(proto_words)
    for i, src in enumerate(proto_words):
        w[off_p + i * L.MG_PROTO_WORDS : off_p + (i + 1) * L.MG_PROTO_WORDS] = w[src : src + L.MG_PROTO_WORDS]

    off_g = blob.alloc(max(1, len(group_recs)) * L.MG_GRP_WORDS)
    w[L.MGC_OFF_GROUPS] = off_g
    for i, src in enumerate(group_recs):
        w[off_g + i * L.MG_GRP_WORDS : off_g + (i + 1) * L.MG_GRP_WORDS] = w[src : src + L.MG_GRP_WORDS]

    for rel, cell_name in pending_object_rewards:  # reward terms on objects.<cell name>: a per-map game stat slot
        code = cell_codes.get(cell_name)
        if code is not None and cell_defs[code - 1]["kind"] != L.MG_KIND_AGENT:
            w[off_g + rel] = L.MG_GS_OBJECTS + code

    off_i = blob.alloc(max(1, len(invcfg_offsets)) * L.MG_INVCFG_WORDS)
    w[L.MGC_OFF_INVCFG] = off_i
    w[L.MGC_NUM_INVCFG] = len(invcfg_offsets)
    for i, src in enumerate(invcfg_offsets):
        w[off_i + i * L.MG_INVCFG_WORDS : off_i + (i + 1) * L.MG_INVCFG_WORDS] = w[src : src + L.MG_INVCFG_WORDS]

    off_a = blob.alloc(max(1, len(asm_recs)) * L.MG_ASM_WORDS)
    w[L.MGC_OFF_ASM_TYPES] = off_a
    w[L.MGC_NUM_ASM_TYPES] = len(asm_recs)
    for i, a in enumerate(asm_recs):
        o = off_a + i * L.MG_ASM_WORDS
        w[o + L.MG_ASM_FIRST_PROTOCOL] = a["first"]
        w[o + L.MG_ASM_NUM_PROTOCOLS] = a["n"]
        w[o + L.MG_ASM_ALLOW_PARTIAL] = a["partial"]
        w[o + L.MG_ASM_MAX_USES] = _i32(a["max_uses"])
        w[o + L.MG_ASM_CLIP_IMMUNE] = a["immune"]
        w[o + L.MG_ASM_START_CLIPPED] = a["start_clipped"]
        w[o + L.MG_ASM_CHEST_DISTANCE] = a["chest_distance"]
        w[o + L.MG_ASM_VIBE_DEPENDENT] = a["dependent"]
        w[o + L.MG_ASM_IDLE_PROTOCOL] = a["idle"]
    w[L.MGC_HAS_VIBE_ASM] = int(any(a["dependent"] for a in asm_recs) or game.clipper is not None)

    off_ch = blob.alloc(max(1, len(chest_recs)) * L.MG_CHEST_WORDS)
    w[L.MGC_OFF_CHEST_TYPES] = off_ch
    w[L.MGC_NUM_CHEST_TYPES] = len(chest_recs)
    for i, c in enumerate(chest_recs):
        o = off_ch + i * L.MG_CHEST_WORDS
        w[o + L.MG_CHEST_INVCFG] = c["invcfg"]
        blob.put_list(o + L.MG_CHEST_INITIAL_INV, c["initial"])
        w[o + L.MG_CHEST_NUM_VIBES] = len(c["vt"])
        for j, (v, deltas) in enumerate(c["vt"]):
            w[o + L.MG_CHEST_VIBE_IDS + j] = v
            blob.put_list(o + L.MG_CHEST_VIBE_LISTS + j * L.MG_LIST_WORDS, deltas)

    off_ct = blob.alloc((len(cell_defs) + 1) * L.MG_CELL_WORDS)  # record 0 = empty
    w[L.MGC_OFF_CELLTYPES] = off_ct
    w[L.MGC_NUM_CELLTYPES] = len(cell_defs) + 1
    for i, c in enumerate(cell_defs, start=1):
        o = off_ct + i * L.MG_CELL_WORDS
        w[o + L.MG_CELL_KIND] = c["kind"]
        w[o + L.MG_CELL_TYPE_ID] = c["type_id"]
        w[o + L.MG_CELL_PARAM] = c["param"]
        w[o + L.MG_CELL_VIBE] = c["vibe"]
        w[o + L.MG_CELL_NUM_TAGS] = len(c["tags"])
        for j, t in enumerate(c["tags"]):
            w[o + L.MG_CELL_TAGS + j] = t

    off_f = blob.alloc(L.MG_RES_STRIDE * L.MG_MAX_INV_TOKENS)
    w[L.MGC_OFF_INV_FEATURES] = off_f
    for r, name in enumerate(resource_names):
        w[off_f + r * L.MG_MAX_INV_TOKENS] = fid[f"inv:{name}"]
        for p in range(1, ntok):
            w[off_f + r * L.MG_MAX_INV_TOKENS + p] = fid[f"inv:{name}:p{p}"]
    off_pf = blob.alloc(2 * L.MG_RES_STRIDE)
    w[L.MGC_OFF_PROTO_FEATURES] = off_pf
    if game.protocol_details_obs:
        for r, name in enumerate(resource_names):
            w[off_pf + r] = fid[f"protocol_input:{name}"]
            w[off_pf + L.MG_RES_STRIDE + r] = fid[f"protocol_output:{name}"]

    w[L.MGC_TOTAL_WORDS] = len(w)
    words = np.asarray([_i32(x) for x in w], dtype=np.int32)
    type_names = {0: "agent"}
    type_names.update({v: k for k, v in type_id.items()})
    return LoweredConfig(
        words=words,
        cell_codes=cell_codes,
        action_names=action_names,
        resource_names=resource_names,
        vibe_names=vibe_names,
        tag_names=tag_names,
        feature_ids=fid,
        feature_normalizations={i: norm for i, _, norm in feats},
        type_names=type_names,
        group_names=group_names,
        cell_kind={i: c["kind"] for i, c in enumerate(cell_defs, start=1)},
        cell_names={i: c["name"] for i, c in enumerate(cell_defs, start=1)},
        num_agents=game.num_agents,
    )
