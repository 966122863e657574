"""MettaGrid::grid_objects() (bindings/mettagrid_c.cpp:754-868) for one env of a batch: the dicts that
`Simulation`, the reference's scripted policies and its tests read, built on the host from the raw
records of the C-ABI (mg_read). A debugging / tooling path, like in the reference: nothing here is on
the step path.

Object ids are the reference's: creation order, row-major over the map's non-empty cells, from 1
(core/grid.hpp:56-68). Feature entries (`obj[feature name] = value`, :800-806) come from the object's
token record, which is its obs_features() list; inventories keep the iteration order of the
reference's unordered_map.
"""

from __future__ import annotations

import numpy as np

from .batch import Batch, order_list
from .layout import L
from .lowering import LoweredConfig


def _feature_names(lowered: LoweredConfig) -> dict[int, str]:
    return {i: n for n, i in lowered.feature_ids.items()}


def _tokens(rec: np.ndarray) -> list[tuple[int, int]]:
    n = int(rec[0])
    tok = rec[8 : 8 + 2 * n].view("<u2")
    return [(int(t & 0xFF), int(t >> 8)) for t in tok]


def _inventory(amt: np.ndarray, order: int) -> dict[int, int]:
    return {item: int(amt[item]) for item in order_list(order)}


def _protocol_lists(w: np.ndarray, p: int):
    off = int(w[L.MGC_OFF_PROTOCOLS]) + p * L.MG_PROTO_WORDS

    def lst(o):
        n = int(w[o + L.MG_LIST_N])
        return {int(w[o + 1 + 2 * i]): int(w[o + 2 + 2 * i]) & 0xFFFF for i in range(n)}

    return lst(off + L.MG_PROTO_INPUTS), lst(off + L.MG_PROTO_OUTPUTS), int(w[off + L.MG_PROTO_COOLDOWN]) & 0xFFFF


def _find_protocol(w: np.ndarray, first: int, n: int, key: int, num_agents: int) -> int:
    for i in range(first, first + n):
        off = int(w[L.MGC_OFF_PROTOCOLS]) + i * L.MG_PROTO_WORDS
        k = ((int(w[off + L.MG_PROTO_KEY_HI]) & 0xFFFFFFFF) << 32) | (int(w[off + L.MG_PROTO_KEY_LO]) & 0xFFFFFFFF)
        if k == key and int(w[off + L.MG_PROTO_MIN_AGENTS]) <= num_agents:
            return i
    return -1


def grid_objects(batch: Batch, env: int = 0, min_row: int = -1, max_row: int = -1, min_col: int = -1, max_col: int = -1,
                 ignore_types: list[str] | tuple[str, ...] = ()) -> dict[int, dict]:
    low = batch.lowered
    w = batch.words
    rec_all, lay = batch.read_token_records()
    rec = rec_all[env]
    er = batch.read_env()[env]
    agents = batch.read_agents()[env]
    asms = batch.read_assemblers()[env]
    chests = batch.read_chests()[env]
    codes = batch.pool.codes[int(er["map_idx"])]
    H, W = codes.shape
    fname = _feature_names(low)
    step = int(er["current_step"])
    use_bounds = min_row >= 0 and max_row >= 0 and min_col >= 0 and max_col >= 0
    cell_off = int(w[L.MGC_OFF_CELLTYPES])

    def cell(code, field):
        return int(w[cell_off + code * L.MG_CELL_WORDS + field])

    agent_at = {(int(a["r"]), int(a["c"])): k for k, a in enumerate(agents)}
    out: dict[int, dict] = {}
    obj_id = 0
    k_agent = k_asm = k_chest = 0
    for r0 in range(H):
        for c0 in range(W):
            code = int(codes[r0, c0])
            if code == 0:
                continue
            obj_id += 1
            kind = cell(code, L.MG_CELL_KIND)
            type_id = cell(code, L.MG_CELL_TYPE_ID)
            d: dict = {}
            r, c = r0, c0
            if kind == L.MG_KIND_AGENT:
                k = k_agent
                k_agent += 1
                a = agents[k]
                r, c = int(a["r"]), int(a["c"])
                slot = lay["agents"] + k
            elif kind == L.MG_KIND_ASSEMBLER:
                k = k_asm
                k_asm += 1
                slot = lay["assemblers"] + k
            elif kind == L.MG_KIND_CHEST:
                k = k_chest
                k_chest += 1
                slot = lay["chests"] + k
            else:
                slot = code
            type_name = low.type_names.get(type_id, "agent" if kind == L.MG_KIND_AGENT else "")
            if type_name in ignore_types:
                continue
            if use_bounds and (r < min_row or r >= max_row or c < min_col or c >= max_col):
                continue
            d["id"] = obj_id
            d["type_name"] = type_name
            d["location"] = (c, r)
            d["r"], d["c"] = r, c
            tokens = _tokens(rec[slot])
            if kind == L.MG_KIND_ASSEMBLER:  # observers prepend cooldown_remaining while it is > 0 (assembler.hpp:533-537)
                rem = max(0, int(asms[k]["cooldown_end"]) - step)
                if rem > 0:
                    tokens = [(int(w[L.MGC_FEAT_COOLDOWN]), min(rem, 255))] + tokens
            for f, v in tokens:
                if f in fname:
                    d[fname[f]] = v
            if kind == L.MG_KIND_AGENT:
                a = agents[k]
                grp = int(a["group"])
                goff = int(w[L.MGC_OFF_GROUPS]) + grp * L.MG_GRP_WORDS
                d["inventory"] = _inventory(a["amt"], int(a["order"]))
                d["group_id"] = int(w[goff + L.MG_GRP_GROUP_ID])
                d["group_name"] = low.group_names[grp]
                d["is_frozen"] = bool(a["frozen"] != 0)
                d["freeze_remaining"] = int(a["frozen"])
                d["freeze_duration"] = int(w[goff + L.MG_GRP_FREEZE_DURATION])
                d["vibe"] = int(a["vibe"])
                d["agent_id"] = k
                d["current_stat_reward"] = float(a["current_stat_reward"])
                d["steps_without_motion"] = int(a["steps_without_motion"])
            elif kind == L.MG_KIND_CHEST:
                ch = chests[k]
                d["inventory"] = _inventory(ch["amt"], int(ch["order"]))
            elif kind == L.MG_KIND_ASSEMBLER:
                s = asms[k]
                toff = int(w[L.MGC_OFF_ASM_TYPES]) + cell(code, L.MG_CELL_PARAM) * L.MG_ASM_WORDS
                clipped = bool(int(s["flags"]) & 1)
                d["cooldown_remaining"] = max(0, int(s["cooldown_end"]) - step)
                d["cooldown_duration"] = int(s["cooldown_duration"])
                d["is_clipped"] = clipped
                d["is_clip_immune"] = bool(w[toff + L.MG_ASM_CLIP_IMMUNE])
                d["uses_count"] = int(s["uses_count"])
                d["max_uses"] = int(w[toff + L.MG_ASM_MAX_USES])
                d["allow_partial_usage"] = bool(w[toff + L.MG_ASM_ALLOW_PARTIAL])
                # get_local_vibe (assembler.hpp:353-375): sorted non-zero vibes of the agents around, packed
                vibes, n_around = [], 0
                for dr in (-1, 0, 1):
                    for dc in (-1, 0, 1):
                        kk = agent_at.get((r + dr, c + dc)) if (dr or dc) else None
                        if kk is not None:
                            n_around += 1
                            if int(agents[kk]["vibe"]):
                                vibes.append(int(agents[kk]["vibe"]))
                key = 0
                for v in sorted(vibes):
                    key = ((key << 8) | v) & 0xFFFFFFFFFFFFFFFF
                as_int = key & 0xFFFFFFFF
                d["current_protocol_id"] = as_int - (1 << 32) if as_int & 0x80000000 else as_int
                if clipped:
                    first, n = int(s["unclip_protocol"]), 1 if int(s["unclip_protocol"]) >= 0 else 0
                else:
                    first, n = int(w[toff + L.MG_ASM_FIRST_PROTOCOL]), int(w[toff + L.MG_ASM_NUM_PROTOCOLS])
                p = _find_protocol(w, first, n, key, n_around) if n else -1
                if p < 0 and key != 0 and n:
                    p = _find_protocol(w, first, n, 0, n_around)
                if p >= 0:
                    ins, outs, cd = _protocol_lists(w, p)
                    d["current_protocol_inputs"] = ins
                    d["current_protocol_outputs"] = outs
                    d["current_protocol_cooldown"] = cd
                protos = []  # every protocol the assembler could run now (the reference lists them in hash-map order)
                for pi in range(first, first + n):
                    off = int(w[L.MGC_OFF_PROTOCOLS]) + pi * L.MG_PROTO_WORDS
                    ins, outs, cd = _protocol_lists(w, pi)
                    nv = int(w[off + L.MG_PROTO_NUM_VIBES])
                    protos.append({"inputs": ins, "outputs": outs, "cooldown": cd, "min_agents": int(w[off + L.MG_PROTO_MIN_AGENTS]),
                                   "vibes": [int(w[off + L.MG_PROTO_VIBES + i]) for i in range(nv)]})
                d["protocols"] = protos
            if kind == L.MG_KIND_CHEST:
                coff = int(w[L.MGC_OFF_CHEST_TYPES]) + cell(code, L.MG_CELL_PARAM) * L.MG_CHEST_WORDS
                vt = {}
                for j in range(int(w[coff + L.MG_CHEST_NUM_VIBES])):
                    lo = coff + L.MG_CHEST_VIBE_LISTS + j * L.MG_LIST_WORDS
                    vt[int(w[coff + L.MG_CHEST_VIBE_IDS + j])] = {int(w[lo + 1 + 2 * i]): int(w[lo + 2 + 2 * i]) for i in range(int(w[lo]))}
                d["vibe_transfers"] = vt
            out[obj_id] = d
    return out
