"""Thin object over the C-ABI handle: one batch of environments on one GPU.

All buffers are raw addresses (``int``); `metta_b200.vecenv` passes ``tensor.data_ptr()`` of CUDA
tensors. State readers return numpy views with the record dtypes of csrc/mg_state.h so that the
conformance harness can compare positions, inventories, cooldowns and stats with the oracle.
"""

from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import capi
from .layout import L
from .lowering import LoweredConfig
from .maps import MapPool

R = L.MG_MAX_RESOURCES

AGENT_DTYPE = np.dtype(
    [
        ("r", "<u2"),
        ("c", "<u2"),
        ("prev_r", "<u2"),
        ("prev_c", "<u2"),
        ("vibe", "u1"),
        ("group", "u1"),
        ("frozen", "<i2"),
        ("steps_without_motion", "<u4"),
        ("order", "<u8"),
        ("amt", "<u2", (R,)),
        ("pad0", "<u2"),
        ("current_stat_reward", "<f4"),
        ("episode_reward", "<f4"),
        ("pad1", "<u4"),
    ]
)
ASM_DTYPE = np.dtype(
    [
        ("r", "<u2"),
        ("c", "<u2"),
        ("cell", "u1"),
        ("flags", "u1"),
        ("unclip_protocol", "<i2"),
        ("cooldown_end", "<u4"),
        ("cooldown_duration", "<u4"),
        ("uses_count", "<u4"),
        ("infection_weight", "<u4"),
        ("cache_protocol", "<i4"),
        ("cache_step", "<u4"),
    ]
)
CHEST_DTYPE = np.dtype(
    [("r", "<u2"), ("c", "<u2"), ("cell", "u1"), ("pad", "u1", (3,)), ("order", "<u8"), ("amt", "<u2", (R,)), ("pad1", "<u2", (3,))]
)
ENV_DTYPE = np.dtype(
    [
        ("current_step", "<u4"),
        ("rng_idx", "<u4"),
        ("clip_rng_idx", "<u4"),
        ("seed", "<u4"),
        ("map_idx", "<i4"),
        ("n_asm", "<i4"),
        ("n_chest", "<i4"),
        ("clip_length_scale", "<i4"),
        ("early_reset_step", "<u4"),
        ("episodes", "<u4"),
        ("done", "u1"),
        ("pad", "u1", (3,)),
        ("spare", "<u4", (5,)),
    ]
)
assert AGENT_DTYPE.itemsize == 64 and ASM_DTYPE.itemsize == 32 and CHEST_DTYPE.itemsize == 48 and ENV_DTYPE.itemsize == 64

ST_ENV, ST_GRID, ST_AGENTS, ST_ASMS, ST_CHESTS, ST_AGENT_STATS, ST_GAME_STATS = range(7)
ST_LAST_ACTION, ST_ACTION_SUCCESS, ST_EPISODE_ACCUM, ST_RNG = 7, 8, 9, 10
ST_OBS, ST_REWARDS, ST_TERMINALS, ST_TRUNCATIONS, ST_ACTIONS = 11, 12, 13, 14, 15


def order_list(order: int) -> list[int]:
    """Decode the 4-bit inventory order list of AgentRec/ChestRec."""
    out = []
    order = int(order)
    for _ in range(16):
        nib = order & 0xF
        if nib == 0xF:
            break
        out.append(nib)
        order >>= 4
    return out


class Batch:
    def __init__(self, lowered: LoweredConfig, pool: MapPool, num_envs: int, device: int = 0, clib=None):
        self._c = clib if clib is not None else capi.lib()
        self.lowered = lowered
        self.pool = pool
        self.num_envs = int(num_envs)
        self.num_agents = lowered.num_agents
        self.num_tokens = int(lowered.words[L.MGC_NUM_TOKENS])
        words = lowered.words.copy()
        if words[L.MGC_MAP_H] != pool.codes.shape[1] or words[L.MGC_MAP_W] != pool.codes.shape[2]:
            raise ValueError("map pool dimensions do not match the lowered config")
        words[L.MGC_MAX_ASSEMBLERS] = pool.max_assemblers
        words[L.MGC_MAX_CHESTS] = pool.max_chests
        self.words = np.ascontiguousarray(words, dtype=np.int32)
        h = ctypes.c_void_p()
        self._check(self._c.mg_create(self.words.ctypes.data, len(self.words), self.num_envs, device, ctypes.byref(h)))
        self._h = h
        grids = np.ascontiguousarray(pool.grids, dtype=np.uint16)
        agents = np.ascontiguousarray(pool.agents, dtype=np.int32)
        n_asm = np.ascontiguousarray(pool.n_asm, dtype=np.int32)
        asms = np.ascontiguousarray(pool.asms, dtype=np.int32)
        n_chest = np.ascontiguousarray(pool.n_chest, dtype=np.int32)
        chests = np.ascontiguousarray(pool.chests, dtype=np.int32)
        self._check(
            self._c.mg_load_maps(
                self._h, len(grids), grids.ctypes.data, agents.ctypes.data, n_asm.ctypes.data, asms.ctypes.data,
                n_chest.ctypes.data, chests.ctypes.data,
            )
        )

    def _check(self, status: int) -> None:
        if status != 0:
            msg = self._c.mg_last_error()
            raise capi.MgError(msg.decode() if msg else f"mettagrid_b200 error {status}")

    # ---- the reference's MettaGrid surface, batched (mettagrid_c.cpp:1007-1041)
    def set_buffers(self, obs: int, terminals: int, truncations: int, rewards: int, actions: int) -> None:
        self._check(self._c.mg_set_buffers(self._h, obs, terminals, truncations, rewards, actions))

    def use_internal_buffers(self) -> None:
        self._check(self._c.mg_set_buffers(self._h, None, None, None, None, None))

    def set_obs_aux(self, row_lengths: int = 0, max_flip: int = 0, obs_mask: int = 0) -> None:
        """Device addresses (0 = off) of i32[N*A] row lengths, an i32 maximum of them and a u8[N*A][T] padding
        mask that every observation pass fills: what ObsShimTokens.forward recomputes per batch."""
        self._check(self._c.mg_set_obs_aux(self._h, row_lengths or None, max_flip or None, obs_mask or None))

    def reset(self, seeds, map_idx, mask=None) -> None:
        seeds = np.ascontiguousarray(seeds, dtype=np.uint32)
        map_idx = np.ascontiguousarray(map_idx, dtype=np.int32)
        assert seeds.shape == (self.num_envs,) and map_idx.shape == (self.num_envs,)
        m = None
        if mask is not None:
            m = np.ascontiguousarray(mask, dtype=np.uint8)
            assert m.shape == (self.num_envs,)
        self._check(self._c.mg_reset(self._h, seeds.ctypes.data, map_idx.ctypes.data, None if m is None else m.ctypes.data))

    def step(self, stream: int = 0) -> None:
        self._check(self._c.mg_step(self._h, stream))

    def step_host(self, actions, obs, rewards, terminals, truncations) -> None:
        self._check(
            self._c.mg_step_host(
                self._h, actions.ctypes.data, obs.ctypes.data, rewards.ctypes.data, terminals.ctypes.data,
                truncations.ctypes.data,
            )
        )

    def step_host_begin(self, actions, obs, rewards, terminals, truncations) -> None:
        """step_host without the wait: returns once the copies and the step are enqueued. The arrays must
        stay untouched until step_host_end()."""
        self._check(
            self._c.mg_step_host_begin(
                self._h, actions.ctypes.data, obs.ctypes.data, rewards.ctypes.data, terminals.ctypes.data,
                truncations.ctypes.data,
            )
        )

    def step_host_end(self) -> None:
        self._check(self._c.mg_step_host_end(self._h))

    def set_host_export(self, changed_rows_only: bool) -> None:
        """step_host with page-locked buffers: False = copy the whole observation buffer every call,
        True = after one full copy only write the part of each row that changed (the caller must not
        write to that buffer between calls)."""
        self._check(self._c.mg_set_host_export(self._h, 1 if changed_rows_only else 0))

    def transfer_bytes(self) -> tuple[int, int]:
        """(host->device, device->host) bytes step_host has moved so far."""
        v = np.zeros(2, dtype=np.uint64)
        self._check(self._c.mg_transfer_bytes(self._h, v.ctypes.data, v.ctypes.data + 8))
        return int(v[0]), int(v[1])

    def profile_step(self, stream: int = 0) -> tuple[float, float]:
        ms = np.zeros(2, dtype=np.float32)
        self._check(self._c.mg_profile_step(self._h, stream, ms.ctypes.data))
        return float(ms[0]), float(ms[1])

    def compute_observations(self, stream: int = 0) -> None:
        self._check(self._c.mg_compute_observations(self._h, stream))

    def peek_observations(self, obs_out: int, stream: int = 0) -> None:
        """Observations of the current state into a caller-owned device buffer (address `obs_out`,
        u8[N*A][T][3]) through the allocation-free entry point: for code that may not allocate GPU
        memory mid-rollout (monitors, replay writers). At most 262144 agents; calls sharing a GPU
        must not overlap; make the first call on a device before any stream capture."""
        self._check(self._c.mg_peek_observations_noalloc(self._h, obs_out, stream))

    def set_options(self, auto_reset: bool = True, map_stride: int = 0) -> None:
        self._check(self._c.mg_set_options(self._h, int(auto_reset), int(map_stride)))

    def set_early_reset(self, steps) -> None:
        s = np.ascontiguousarray(steps, dtype=np.uint32)
        assert s.shape == (self.num_envs,)
        self._check(self._c.mg_set_early_reset(self._h, s.ctypes.data))

    def set_inventory(self, env: int, agent: int, inventory: dict[int, int]) -> None:
        """MettaGrid::set_inventory (mettagrid_c.cpp:1007-1041). The reference receives the dict as an
        unordered_map and applies it in that map's iteration order (agent.cpp:68-71), which decides the
        order of the inventory tokens afterwards; the same order is computed here."""
        from .lowering import map_order

        n = len(inventory)
        pairs = map_order([(int(k), int(v)) for k, v in inventory.items()])
        buf = np.asarray([p[0] for p in pairs] + [p[1] for p in pairs], dtype=np.int32)
        self._check(self._c.mg_set_inventory(self._h, env, agent, buf.ctypes.data if n else None, n))

    def sync(self) -> None:
        self._check(self._c.mg_sync(self._h))

    @property
    def kernel_launches(self) -> int:
        return int(self._c.mg_kernel_launches(self._h))

    # ---- state readers
    def _read(self, what: int, dtype, shape) -> np.ndarray:
        nbytes = int(self._c.mg_sizeof(self._h, what))
        if nbytes < 0:
            raise capi.MgError("unknown state array")
        out = np.zeros(nbytes, dtype=np.uint8)
        self._check(self._c.mg_read(self._h, what, out.ctypes.data, nbytes))
        return out.view(dtype).reshape(shape)

    def _write(self, what: int, arr: np.ndarray) -> None:
        a = np.ascontiguousarray(arr)
        self._check(self._c.mg_write(self._h, what, a.ctypes.data, a.nbytes))

    def read_env(self) -> np.ndarray:
        return self._read(ST_ENV, ENV_DTYPE, (self.num_envs,))

    def read_agents(self) -> np.ndarray:
        return self._read(ST_AGENTS, AGENT_DTYPE, (self.num_envs, self.num_agents))

    def read_assemblers(self) -> np.ndarray:
        return self._read(ST_ASMS, ASM_DTYPE, (self.num_envs, -1))

    def read_chests(self) -> np.ndarray:
        return self._read(ST_CHESTS, CHEST_DTYPE, (self.num_envs, -1))

    def read_grid(self) -> np.ndarray:
        H, W = int(self.words[L.MGC_MAP_H]), int(self.words[L.MGC_MAP_W])
        g = self._read(ST_GRID, np.uint16, (self.num_envs, -1))
        return g[:, : H * W].reshape(self.num_envs, H, W)

    def read_agent_stats(self) -> np.ndarray:
        return self._read(ST_AGENT_STATS, np.float32, (self.num_envs, self.num_agents, L.MG_AS_COUNT))

    def read_game_stats(self) -> np.ndarray:
        return self._read(ST_GAME_STATS, np.float32, (self.num_envs, L.MG_GS_COUNT))

    def read_action_success(self) -> np.ndarray:
        return self._read(ST_ACTION_SUCCESS, np.uint8, (self.num_envs, self.num_agents)).astype(bool)

    def read_token_records(self) -> tuple[np.ndarray, dict]:
        """(u8 [N, records per env, bytes per record], layout): what every object currently shows to observers."""
        lay = np.zeros(5, dtype=np.int32)
        self._check(self._c.mg_token_record_layout(self._h, lay.ctypes.data))
        rec = self._read(18, np.uint8, (self.num_envs, int(lay[1]), int(lay[0])))
        return rec, {"agents": int(lay[2]), "assemblers": int(lay[3]), "chests": int(lay[4])}

    def read_invalid_ids(self) -> np.ndarray:
        """i32 [N, A, MG_INVALID_WORDS]: per agent the out-of-range action ids it sent this episode as (id, count)
        pairs, the count of any further ones, and the number of pairs in use (action.invalid_index.<id>)."""
        return self._read(17, np.int32, (self.num_envs, self.num_agents, L.MG_INVALID_WORDS))

    def read_episode_rewards(self) -> np.ndarray:
        return self.read_agents()["episode_reward"].copy()

    def read_episode_accum(self) -> np.ndarray:
        return self._read(ST_EPISODE_ACCUM, np.float64, (-1,))

    def read_rng(self) -> np.ndarray:
        return self._read(ST_RNG, np.uint32, (self.num_envs, 624))

    def read_obs(self) -> np.ndarray:
        """The bound observation buffer, copied to the host: u8 [N*A, T, 3]."""
        return self._read(ST_OBS, np.uint8, (self.num_envs * self.num_agents, -1, 3))

    def read_buffer(self, what: int, dtype, shape) -> np.ndarray:
        return self._read(what, dtype, shape)

    def write_actions(self, actions) -> None:
        self._write(ST_ACTIONS, np.ascontiguousarray(actions, dtype=np.int32).reshape(-1))

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._c.mg_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def bind_library(clib) -> None:
    """Declare argument types on a CDLL exporting the C-ABI (used for the product library)."""
    capi._declare(clib)


__all__ = ["Batch", "order_list", "AGENT_DTYPE", "ASM_DTYPE", "CHEST_DTYPE", "ENV_DTYPE", "Optional"]
