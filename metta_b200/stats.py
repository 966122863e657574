"""Dense stat slots -> the reference's stat names.

The reference keeps string-keyed float maps (systems/stats_tracker.hpp:14-65) and reports them
through get_episode_stats() (bindings/mettagrid_c.cpp:936-953); the simulator keeps dense float
slots (include/mg_config.h MG_AS_*/MG_GS_*). These tables translate back, omitting slots that are
zero exactly like a never-touched map key is absent. Names: SURVEY.md Appendix C.
"""

from __future__ import annotations

import numpy as np

from .layout import L
from .lowering import LoweredConfig

ACTION_KINDS = ["noop", "move", "change_vibe"]


def agent_stat_names(lowered: LoweredConfig, group: int) -> dict[int, list[str]]:
    """slot -> names it stands for, for an agent of agent-config `group`."""
    g = lowered.group_names[group]
    out: dict[int, list[str]] = {}
    for k, kind in enumerate(ACTION_KINDS):
        out[L.MG_AS_ACTION_SUCCESS + k] = [f"action.{kind}.success"]
        out[L.MG_AS_ACTION_FAILED + k] = [f"action.{kind}.failed"]
    out[L.MG_AS_ACTION_FAILED_ANY] = ["action.failed"]
    out[L.MG_AS_FROZEN_TICKS] = ["status.frozen.ticks", f"status.frozen.ticks.{g}"]
    out[L.MG_AS_MAX_STEPS_WITHOUT_MOTION] = ["status.max_steps_without_motion"]
    out[L.MG_AS_SWAP] = ["actions.swap"]
    out[L.MG_AS_DIVERSITY] = ["inventory.diversity"]
    out[L.MG_AS_DAMAGE_ITEMS_LOST] = ["damage.items_lost"]
    out[L.MG_AS_DAMAGE_TRIGGERED] = ["damage.triggered"]
    out[L.MG_AS_INVALID_INDEX] = ["action.invalid_index"]
    out[L.MG_AS_ATTACK_FRIENDLY_FIRE] = [f"action.attack.{g}.friendly_fire"]
    out[L.MG_AS_TRANSFER_COUNT] = [f"action.transfer.{g}.count"]
    for j, other in enumerate(lowered.group_names):
        out[L.MG_AS_ATTACK_BLOCKED_BY + j] = [f"action.attack.{g}.blocked_by.{other}"]
        out[L.MG_AS_ATTACK_HIT + j] = [f"action.attack.{g}.hit.{other}"]
        out[L.MG_AS_ATTACK_HIT_BY + j] = [f"action.attack.{g}.hit_by.{other}"]
    for r, res in enumerate(lowered.resource_names):
        out[L.MG_AS_AMOUNT + r] = [f"{res}.amount"]
        out[L.MG_AS_GAINED + r] = [f"{res}.gained"]
        out[L.MG_AS_LOST + r] = [f"{res}.lost"]
        out[L.MG_AS_CHEST_DEPOSITED + r] = [f"chest.{res}.deposited_by_agent"]
        out[L.MG_AS_CHEST_WITHDRAWN + r] = [f"chest.{res}.withdrawn_by_agent"]
        out[L.MG_AS_DAMAGED + r] = [f"damaged.{res}"]
        out[L.MG_AS_TRANSFER_SELF + r] = [f"action.transfer.{g}.{res}.self"]
        for j, other in enumerate(lowered.group_names):
            out[L.MG_AS_TRANSFER_TO + j * L.MG_RES_STRIDE + r] = [f"action.transfer.{g}.{res}.to.{other}"]
    return out


def game_stat_names(lowered: LoweredConfig) -> dict[int, str]:
    out = {
        L.MG_GS_TOKENS_WRITTEN: "tokens_written",
        L.MG_GS_TOKENS_DROPPED: "tokens_dropped",
        L.MG_GS_TOKENS_FREE_SPACE: "tokens_free_space",
    }
    for r, res in enumerate(lowered.resource_names):
        out[L.MG_GS_ASM_CREATED + r] = f"assembler.{res}.created"
        out[L.MG_GS_CHEST_DEPOSITED + r] = f"chest.{res}.deposited"
        out[L.MG_GS_CHEST_WITHDRAWN + r] = f"chest.{res}.withdrawn"
        out[L.MG_GS_CHEST_AMOUNT + r] = f"chest.{res}.amount"
    return out


def agent_stats_dict(lowered: LoweredConfig, group: int, slots: np.ndarray, invalid_ids: np.ndarray | None = None) -> dict[str, float]:
    """`invalid_ids`: the agent's i32[MG_INVALID_WORDS] record (Batch.read_invalid_ids) for the per-id breakdown
    action.invalid_index.<id> (mettagrid_c.cpp:565-569); ids beyond the first MG_INVALID_IDS distinct ones of an
    episode are reported together as action.invalid_index.other."""
    names = agent_stat_names(lowered, group)
    out: dict[str, float] = {}
    for slot in np.nonzero(slots)[0]:
        for n in names.get(int(slot), []):
            out[n] = float(slots[slot])
    if invalid_ids is not None:
        n = int(invalid_ids[2 * L.MG_INVALID_IDS + 1])
        for q in range(n):
            out[f"action.invalid_index.{int(invalid_ids[2 * q])}"] = float(invalid_ids[2 * q + 1])
        if invalid_ids[2 * L.MG_INVALID_IDS]:
            out["action.invalid_index.other"] = float(invalid_ids[2 * L.MG_INVALID_IDS])
    return out


def game_stats_dict(lowered: LoweredConfig, slots: np.ndarray, map_codes: np.ndarray | None = None) -> dict[str, float]:
    names = game_stat_names(lowered)
    out = {names[int(s)]: float(slots[s]) for s in np.nonzero(slots)[0] if int(s) in names}
    if len(slots) >= L.MG_GS_OBJECTS + L.MG_MAX_CELLTYPES:  # objects.<cell name> live in slots of their own
        for code, name in lowered.cell_names.items():
            v = float(slots[L.MG_GS_OBJECTS + code]) if code < L.MG_MAX_CELLTYPES else 0.0
            if v:
                out[f"objects.{name}"] = v
        return out
    if map_codes is not None:  # objects.<cell name>, counted at construction (mettagrid_c.cpp:203,224,235)
        codes, counts = np.unique(map_codes, return_counts=True)
        for code, n in zip(codes, counts):
            code = int(code)
            if code and lowered.cell_kind.get(code) != L.MG_KIND_AGENT:
                out[f"objects.{lowered.cell_names[code]}"] = float(n)
    return out
