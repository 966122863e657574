"""ctypes wrapper of oracle/libmg_oracle.so (mg_oracle.c). TEST INFRASTRUCTURE, not product code."""

from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libmg_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "mg_oracle.c")
    hdr = os.path.join(_HERE, "..", "include", "mg_config.h")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libmg_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        l = ctypes.CDLL(_LIB)
        vp, i32, u32 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_uint32
        l.mgo_create.argtypes = [vp, vp, u32]
        l.mgo_create.restype = vp
        l.mgo_destroy.argtypes = [vp]
        l.mgo_set_buffers.argtypes = [vp] * 6
        l.mgo_step.argtypes = [vp]
        l.mgo_error.argtypes = [vp]
        l.mgo_current_step.argtypes = [vp]
        l.mgo_current_step.restype = u32
        l.mgo_num_assemblers.argtypes = [vp]
        l.mgo_num_chests.argtypes = [vp]
        for name in ("mgo_episode_rewards", "mgo_action_success", "mgo_game_stats"):
            getattr(l, name).argtypes = [vp]
            getattr(l, name).restype = vp
        l.mgo_agent_stats.argtypes = [vp, i32]
        l.mgo_agent_stats.restype = vp
        for name in ("mgo_agent_state", "mgo_assembler_state", "mgo_chest_state"):
            getattr(l, name).argtypes = [vp, i32, vp]
        l.mgo_set_inventory.argtypes = [vp, i32, vp, vp, i32]
        l.mgo_invalid_ids.argtypes = [vp, i32]
        l.mgo_invalid_ids.restype = vp
        l.mgo_get_rng.argtypes = [vp, vp]
        l.mgo_set_rng.argtypes = [vp, vp]
        _lib = l
    return _lib


class OracleEnv:
    """One environment, stepped on the CPU by the C restatement of the reference algorithm."""

    def __init__(self, words: np.ndarray, map_codes: np.ndarray, seed: int):
        from metta_b200.layout import L

        self.L = L
        self.words = np.ascontiguousarray(words, dtype=np.int32)
        codes = np.ascontiguousarray(map_codes, dtype=np.uint8)
        assert codes.shape == (self.words[L.MGC_MAP_H], self.words[L.MGC_MAP_W]), codes.shape
        self._h = lib().mgo_create(self.words.ctypes.data, codes.ctypes.data, seed & 0xFFFFFFFF)
        if not self._h:
            raise RuntimeError("oracle: bad config blob")
        if lib().mgo_error(self._h):
            raise RuntimeError(f"oracle: map/config mismatch (code {lib().mgo_error(self._h)})")
        A, T = int(self.words[L.MGC_NUM_AGENTS]), int(self.words[L.MGC_NUM_TOKENS])
        self.num_agents, self.num_tokens = A, T
        self.observations = np.zeros((A, T, 3), dtype=np.uint8)
        self.terminals = np.zeros(A, dtype=np.bool_)
        self.truncations = np.zeros(A, dtype=np.bool_)
        self.rewards = np.zeros(A, dtype=np.float32)
        self.actions = np.zeros(A, dtype=np.int32)
        lib().mgo_set_buffers(
            self._h,
            self.observations.ctypes.data,
            self.terminals.ctypes.data,
            self.truncations.ctypes.data,
            self.rewards.ctypes.data,
            self.actions.ctypes.data,
        )

    def step(self, actions=None) -> None:
        if actions is not None:
            self.actions[:] = actions
        lib().mgo_step(self._h)
        if lib().mgo_error(self._h):
            raise RuntimeError("oracle: internal error")

    @property
    def current_step(self) -> int:
        return int(lib().mgo_current_step(self._h))

    def _arr(self, ptr, n, dtype):
        return np.ctypeslib.as_array(ctypes.cast(ptr, ctypes.POINTER(dtype)), shape=(n,)).copy()

    def episode_rewards(self) -> np.ndarray:
        return self._arr(lib().mgo_episode_rewards(self._h), self.num_agents, ctypes.c_float)

    def action_success(self) -> np.ndarray:
        return self._arr(lib().mgo_action_success(self._h), self.num_agents, ctypes.c_uint8).astype(bool)

    def game_stats(self) -> np.ndarray:
        return self._arr(lib().mgo_game_stats(self._h), self.L.MG_GS_COUNT, ctypes.c_float)

    def agent_stats(self) -> np.ndarray:
        return np.stack(
            [self._arr(lib().mgo_agent_stats(self._h, i), self.L.MG_AS_COUNT, ctypes.c_float) for i in range(self.num_agents)]
        )

    def agent_state(self) -> np.ndarray:
        out = np.zeros((self.num_agents, 6 + 2 * self.L.MG_RES_STRIDE), dtype=np.int32)
        for i in range(self.num_agents):
            lib().mgo_agent_state(self._h, i, out[i].ctypes.data)
        return out

    def assembler_state(self) -> np.ndarray:
        n = lib().mgo_num_assemblers(self._h)
        out = np.zeros((n, 8), dtype=np.int32)
        for i in range(n):
            lib().mgo_assembler_state(self._h, i, out[i].ctypes.data)
        return out

    def chest_state(self) -> np.ndarray:
        n = lib().mgo_num_chests(self._h)
        out = np.zeros((n, 3 + 2 * self.L.MG_RES_STRIDE), dtype=np.int32)
        for i in range(n):
            lib().mgo_chest_state(self._h, i, out[i].ctypes.data)
        return out

    def set_inventory(self, agent: int, inventory: dict[int, int]) -> None:
        # applied in the iteration order of the unordered_map pybind11 builds from the dict (agent.cpp:68-71)
        from metta_b200.lowering import map_order

        pairs = map_order([(int(k), int(v)) for k, v in inventory.items()])
        items = np.asarray([p[0] for p in pairs], dtype=np.int32)
        amounts = np.asarray([p[1] for p in pairs], dtype=np.int32)
        lib().mgo_set_inventory(self._h, agent, items.ctypes.data, amounts.ctypes.data, len(items))

    def invalid_ids(self) -> np.ndarray:
        """i32 [A][MG_INVALID_WORDS]: the record behind action.invalid_index.<id>."""
        n = self.L.MG_INVALID_WORDS
        return np.stack([self._arr(lib().mgo_invalid_ids(self._h, i), n, ctypes.c_int32) for i in range(self.num_agents)])

    def get_rng(self) -> np.ndarray:
        """The env generator: 624 state words followed by the position (std::mt19937 _M_x, _M_p). Test hook."""
        out = np.zeros(625, dtype=np.uint32)
        lib().mgo_get_rng(self._h, out.ctypes.data)
        return out

    def set_rng(self, state625: np.ndarray) -> None:
        s = np.ascontiguousarray(state625, dtype=np.uint32)
        assert s.shape == (625,)
        lib().mgo_set_rng(self._h, s.ctypes.data)

    def close(self) -> None:
        if self._h:
            lib().mgo_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
