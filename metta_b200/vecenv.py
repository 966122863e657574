"""GpuMettaGridVecEnv: thousands of MettaGrid envs on one B200 behind the reference's env API.

What it stands in for:
  * pufferlib's vec-env contract (packages/pufferlib-core/src/pufferlib/vector.py:121-183):
    ``async_reset / recv / send / close``, ``num_agents``, ``single_observation_space``,
    ``single_action_space``, ``driver_env``;
  * MettaGridPufferEnv's step semantics (mettagrid/envs/mettagrid_puffer_env.py:149-183):
    shared buffers, auto-reset at the start of the step after an episode ended, same seed;
  * metta/rl's TrainingEnvironment (metta/rl/training/training_environment.py:85-119,253-275):
    ``get_observations / send_actions``, with a contiguous env slice.

State stays in HBM; observations, rewards, terminals and truncations are torch CUDA tensors that
the kernels write in place, so a policy can read them without a host round trip. ``host=True``
gives numpy arrays instead (the drop-in path for code that expects the reference's arrays).
There is no CPU implementation: without the CUDA library or a GPU, construction raises.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Optional, Sequence

import time

import numpy as np

from .batch import Batch
from .config import MettaGridConfig
from .layout import L
from .lowering import LoweredConfig, lower_game_config
from .maps import MapPool, build_pool, names_to_codes
from .stats import agent_stat_names, game_stat_names, game_stats_dict


@dataclass
class Box:  # gymnasium.spaces.Box stand-in (gymnasium is not a dependency)
    low: int
    high: int
    shape: tuple
    dtype: Any


@dataclass
class Discrete:
    n: int
    dtype: Any = np.int32


class GpuMettaGridVecEnv:
    def __init__(
        self,
        cfg: MettaGridConfig | Any,
        num_envs: int,
        maps: Optional[Sequence] = None,
        map_codes: Optional[np.ndarray] = None,
        seed: int = 0,
        device: int | str = 0,
        host: bool = False,
        map_stride: int = 0,
        obs_mask: bool = False,
        first_env: int = 0,
    ):
        import torch

        if not torch.cuda.is_available():
            raise RuntimeError("GpuMettaGridVecEnv needs a CUDA device; there is no CPU simulator in this package")
        self.cfg = MettaGridConfig.from_reference(cfg)
        game = self.cfg.game
        if map_codes is None:
            if maps is None:
                raise ValueError("pass `maps` (cell-name grids, as the reference's map builders produce) or `map_codes`")
            first = np.asarray(maps[0])
            lowered = lower_game_config(game, first.shape[0], first.shape[1])
            map_codes = np.stack([names_to_codes(lowered, m) for m in maps])
        else:
            map_codes = np.asarray(map_codes, dtype=np.uint8)
            lowered = lower_game_config(game, map_codes.shape[1], map_codes.shape[2])
        self.lowered: LoweredConfig = lowered
        self.pool: MapPool = build_pool(lowered, map_codes)
        self.num_envs = int(num_envs)
        self.agents_per_env = game.num_agents
        self.num_agents = self.num_envs * self.agents_per_env
        dev = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        self.device = dev
        self._torch = torch
        self.batch = Batch(lowered, self.pool, self.num_envs, device=dev.index or 0)
        self.batch.set_options(auto_reset=True, map_stride=map_stride)
        T = game.obs.num_tokens
        self.single_observation_space = Box(0, 255, (T, 3), np.uint8)
        self.single_action_space = Discrete(len(lowered.action_names))
        self.action_names = list(lowered.action_names)
        self.host = host
        NA = self.num_agents
        self.observations = torch.full((NA, T, 3), 255, dtype=torch.uint8, device=dev)
        self.rewards = torch.zeros(NA, dtype=torch.float32, device=dev)
        self.terminals = torch.zeros(NA, dtype=torch.bool, device=dev)
        self.truncations = torch.zeros(NA, dtype=torch.bool, device=dev)
        self.actions = torch.zeros(NA, dtype=torch.int32, device=dev)
        self.teacher_actions = torch.zeros(NA, dtype=torch.int32, device=dev)
        self.masks = torch.ones(NA, dtype=torch.bool, device=dev)
        # this object's envs are envs [first_env, first_env + num_envs) of a larger job (another rank, or the other
        # half of a GpuMettaGridVecEnvPair): seeds, maps, agent ids and env_index in infos count from there
        self.first_env = int(first_env)
        self.agent_ids = np.arange(NA) + self.first_env * self.agents_per_env
        self.batch.set_buffers(
            self.observations.data_ptr(), self.terminals.data_ptr(), self.truncations.data_ptr(),
            self.rewards.data_ptr(), self.actions.data_ptr(),
        )
        # what the policy's first op would recompute from the observations (obs_shim.py:181-196): tokens per
        # row, their maximum and, on request, the padding mask, written by the observation kernel
        self.row_lengths = torch.zeros(NA, dtype=torch.int32, device=dev)
        self.max_flip = torch.zeros((), dtype=torch.int32, device=dev)
        self.obs_mask = torch.ones((NA, T), dtype=torch.bool, device=dev) if obs_mask else None
        self._want_obs_mask = obs_mask
        self.batch.set_obs_aux(self.row_lengths.data_ptr(), self.max_flip.data_ptr(),
                               self.obs_mask.data_ptr() if obs_mask else 0)
        if host:
            # page-locked arrays: the copy engines and the export kernel reach them directly, and since this
            # object is the only writer of them, only the parts of the rows that changed cross the host link
            pin = lambda shape, dtype: torch.zeros(shape, dtype=dtype, pin_memory=True).numpy()  # noqa: E731
            self._h_obs = pin((NA, T, 3), torch.uint8)
            self._h_obs[...] = 255
            self._h_rew = pin((NA,), torch.float32)
            self._h_term = pin((NA,), torch.uint8).view(np.bool_)
            self._h_trunc = pin((NA,), torch.uint8).view(np.bool_)
            self._h_act = pin((NA,), torch.int32)
            self.batch.set_host_export(True)
        self._inflight = False  # host mode: a step has been enqueued by send() and not yet waited for by recv()
        self._seed = seed
        self._map_stride = map_stride
        self._pending = None
        self._episodes_reported = 0.0
        self._accum_prev: Optional[np.ndarray] = None
        self.driver_env = self
        # Episode ends are decided by step counts alone (max_steps, early reset), so the host knows
        # which envs finish on which step without reading a flag back from the GPU.
        self.collect_infos = True
        self._max_steps = int(game.max_steps)
        self._steps = np.zeros(self.num_envs, dtype=np.int64)
        self._early = np.zeros(self.num_envs, dtype=np.int64)
        self._done = np.zeros(self.num_envs, dtype=bool)
        self._seeds = np.zeros(self.num_envs, dtype=np.int64)
        self._names_a = [agent_stat_names(lowered, g) for g in range(len(lowered.group_names))]

    # ---- pufferlib contract
    def async_reset(self, seed: Optional[int] = None) -> None:
        if seed is not None:
            self._seed = seed
        if self._pending is not None:
            self._apply_pending_config()
        if self._inflight:
            self.batch.step_host_end()
            self._inflight = False
        seeds = (self.first_env + np.arange(self.num_envs, dtype=np.int64) + int(self._seed)) & 0xFFFFFFFF
        maps = (self.first_env + np.arange(self.num_envs)) % len(self.pool.codes)
        self.batch.reset(seeds.astype(np.uint32), maps.astype(np.int32))
        self._accum_prev = None
        self._steps[:] = 0
        self._done[:] = False
        self._early[:] = 0
        self._seeds[:] = seeds
        if self.host:
            self._download()
            self.batch.set_host_export(True)  # the host arrays were just rewritten here: next step copies everything once

    def reset(self, seed: Optional[int] = None):
        self.async_reset(seed)
        return self.recv()[0], {}

    def send(self, actions) -> None:
        torch = self._torch
        # envs that ended on the last step start a new episode at the start of this one
        # (MettaGridPufferEnv.step), then every env advances by one step
        self._steps[self._done] = 0
        self._early[self._done] = 0
        self._steps += 1
        limit = np.where(self._early > 0, self._early, np.iinfo(np.int64).max)
        if self._max_steps > 0:
            limit = np.minimum(limit, self._max_steps)
        self._done = self._steps >= limit
        if self.host:
            # enqueue and return (pufferlib's send); recv() waits. The caller's thread is free in between -- with a
            # second object (GpuMettaGridVecEnvPair) it is acting on the other half of the envs
            if self._inflight:
                self.batch.step_host_end()
            self._h_act[...] = np.asarray(actions).reshape(-1)
            self.batch.step_host_begin(self._h_act, self._h_obs, self._h_rew, self._h_term.view(np.uint8), self._h_trunc.view(np.uint8))
            self._inflight = True
            return
        if isinstance(actions, np.ndarray):
            actions = torch.from_numpy(np.ascontiguousarray(actions.reshape(-1), dtype=np.int32))
        self.actions.copy_(actions.reshape(-1).to(torch.int32), non_blocking=True)
        self.batch.step(torch.cuda.current_stream(self.device).cuda_stream)

    def _wait(self) -> None:
        """Host mode: let the step that send() enqueued finish (everything that reads the batch does this first)."""
        if self._inflight:
            self.batch.step_host_end()
            self._inflight = False

    def recv(self):
        self._wait()
        infos = self._infos()
        if self.host:
            return (self._h_obs, self._h_rew, self._h_term, self._h_trunc, None, infos, self.agent_ids, None)
        return (self.observations, self.rewards, self.terminals, self.truncations, self.teacher_actions, infos,
                self.agent_ids, self.masks)

    def step(self, actions):
        self.send(actions)
        o, r, d, t, _, infos, _, _ = self.recv()
        return o, r, d, t, infos

    def close(self) -> None:
        if self._inflight:
            self.batch.step_host_end()
            self._inflight = False
        self.batch.close()

    # ---- metta/rl TrainingEnvironment surface
    def get_observations(self):
        o, r, d, t, ta, infos, ids, mask = self.recv()
        return o, r, d, t, ta, infos, slice(0, self.num_agents), mask, self.num_agents

    def send_actions(self, actions) -> None:
        self.send(actions)

    # ---- helpers
    def _download(self) -> None:
        self._h_obs[...] = self.observations.cpu().numpy()
        self._h_rew[...] = self.rewards.cpu().numpy()
        self._h_term[...] = self.terminals.cpu().numpy()
        self._h_trunc[...] = self.truncations.cpu().numpy()

    def episode_stats(self) -> dict[str, float]:
        """Sums over all episodes finished on this GPU since construction (a dense vector that
        ``reduce_episode_stats`` all-reduces across ranks)."""
        self._wait()
        acc = self.batch.read_episode_accum()
        names_a = agent_stat_names(self.lowered, 0)
        names_g = game_stat_names(self.lowered)
        out = {"episodes": float(acc[0]), "episode_reward_sum": float(acc[1])}
        S = L.MG_AS_COUNT
        for s in np.nonzero(acc[2 : 2 + S])[0]:
            for n in names_a.get(int(s), []):
                out["agent/" + n] = float(acc[2 + s])
        for s in np.nonzero(acc[2 + S :])[0]:
            if int(s) in names_g:
                out["game/" + names_g[int(s)]] = float(acc[2 + S + s])
        return out

    def set_early_reset(self, steps) -> None:
        """EarlyResetHandler (mettagrid/envs/early_reset_handler.py:12-20): truncate the running episode of
        env i once its step count reaches steps[i] (0 = never)."""
        self._wait()
        steps = np.ascontiguousarray(steps, dtype=np.uint32)
        self.batch.set_early_reset(steps)
        self._early[:] = steps

    # ---- the read-only properties of MettaGrid (mettagrid_c.cpp:1007-1041) that callers ask for
    @property
    def obs_width(self) -> int:
        return int(self.cfg.game.obs.width)

    @property
    def obs_height(self) -> int:
        return int(self.cfg.game.obs.height)

    @property
    def max_steps(self) -> int:
        return self._max_steps

    @property
    def map_width(self) -> int:
        return int(self.pool.codes.shape[2])

    @property
    def map_height(self) -> int:
        return int(self.pool.codes.shape[1])

    @property
    def resource_names(self) -> list[str]:
        return list(self.lowered.resource_names)

    @property
    def object_type_names(self) -> list[str]:
        # MettaGrid::object_type_names: sized by the number of object configs (map cell names) first, then by
        # the largest type id (mettagrid_c.cpp:163-178)
        n = max(len(self.lowered.cell_codes), max(self.lowered.type_names) + 1) if self.lowered.type_names else 0
        return [self.lowered.type_names.get(i, "") for i in range(n)]

    @property
    def current_step(self) -> np.ndarray:
        """Steps taken in the running episode, per env (host mirror, no device read-back)."""
        return self._steps.copy()

    def action_success(self) -> np.ndarray:
        """bool [num_envs, agents_per_env]: MettaGrid::action_success of the last step."""
        self._wait()
        return self.batch.read_action_success()

    def grid_objects(self, env_index: int = 0, min_row: int = -1, max_row: int = -1, min_col: int = -1, max_col: int = -1,
                     ignore_types: Sequence[str] = ()) -> dict[int, dict]:
        """MettaGrid::grid_objects (mettagrid_c.cpp:754-868) of one env: the reference's dicts, same ids, keys and
        inventory iteration order (a host-side tooling path: it reads the env's records back)."""
        self._wait()
        from .grid_objects import grid_objects

        return grid_objects(self.batch, env_index, min_row, max_row, min_col, max_col, tuple(ignore_types))

    def set_mg_config(self, cfg, maps: Optional[Sequence] = None, map_codes: Optional[np.ndarray] = None) -> None:
        """CurriculumEnv's hook (metta/rl/.../curriculum_env.py:105-117): the next `async_reset` builds the
        batch from this config (and these maps; the current pool is kept when none are given and the map
        cell names still resolve)."""
        self._pending = (cfg, maps, map_codes)

    def _apply_pending_config(self) -> None:
        self._wait()
        cfg, maps, map_codes = self._pending
        self._pending = None
        if maps is None and map_codes is None:
            map_codes = self.pool.codes
        self.batch.close()
        self.__init__(cfg, self.num_envs, maps=maps, map_codes=map_codes, seed=self._seed, device=self.device, host=self.host,
                      map_stride=self._map_stride, obs_mask=self._want_obs_mask, first_env=self.first_env)

    def get_episode_rewards(self):
        """f32 [num_envs, agents_per_env] (MettaGrid::get_episode_rewards; CurriculumEnv takes its mean)."""
        self._wait()
        return self.batch.read_episode_rewards()

    def _infos(self) -> list[dict]:
        """One dict per env whose episode ended on the last step, shaped like the reference's
        StatsTracker.on_episode_end (mettagrid/envs/stats_tracker.py:26-76): ``game`` stats, ``agent``
        stats averaged over the env's agents, ``attributes``. The stats are still in HBM at this point:
        the env is rebuilt at the start of the next step."""
        if not self.collect_infos or not self._done.any():
            return []
        idx = np.nonzero(self._done)[0]
        astats = self.batch.read_agent_stats()[idx].astype(np.float64)   # [n, A, S]
        gstats = self.batch.read_game_stats()[idx]
        groups = self.batch.read_agents()["group"][idx]
        rewards = self.batch.read_episode_rewards()[idx]
        recs = self.batch.read_env()[idx]
        now = time.time()
        out = []
        A = self.agents_per_env
        for k, env in enumerate(idx):
            agent: dict[str, float] = {}
            for g in np.unique(groups[k]):
                total = astats[k][groups[k] == g].sum(axis=0)
                names = self._names_a[int(g)]
                for slot in np.nonzero(total)[0]:
                    for n in names.get(int(slot), []):
                        agent[n] = agent.get(n, 0.0) + float(total[slot])
            out.append({
                "game": game_stats_dict(self.lowered, gstats[k], self.pool.codes[int(recs["map_idx"][k])]),
                "agent": {n: v / A for n, v in agent.items()},
                "attributes": {
                    "seed": int(self._seeds[env]), "map_w": int(self.pool.codes.shape[2]), "map_h": int(self.pool.codes.shape[1]),
                    "steps": int(recs["current_step"][k]), "max_steps": self._max_steps, "completion_time": now,
                },
                "episode_rewards": rewards[k].copy(),
                "env_index": self.first_env + int(env),
            })
        return out

    def reduce_episode_stats(self):
        """All-reduce (SUM) the episode accumulators over the process group: the only collective of
        the env path. Returns a float64 tensor on this rank's GPU."""
        self._wait()
        torch = self._torch
        from .sharding import reduce_episode_accumulators

        acc = torch.from_numpy(self.batch.read_episode_accum()).to(self.device)
        return reduce_episode_accumulators(acc)


class GpuMettaGridVecEnvPair:
    """The envs as two halves that are stepped alternately, for numpy callers: `recv()` hands out the halves in
    turn, `send(actions)` answers the half received last and returns at once, so one half's observations cross
    the host link while the caller works on the other half. This is pufferlib's asynchronous contract with
    `batch_size = num_envs / 2`, which the reference's trainer sets up with `async_factor` 2
    (metta/rl/training/training_environment.py:45,152-169, batch.py:10-19): every `recv()` returns the arrays of
    `num_agents / 2` agents and their `agent_ids`.

    Each half is a complete GpuMettaGridVecEnv(host=True) of its own (own handle, own page-locked arrays); seeds,
    maps, agent ids and `env_index` in the infos are those of one env object over all `num_envs`."""

    def __init__(self, cfg, num_envs: int, seed: int = 0, **kw):
        if num_envs % 2:
            raise ValueError("GpuMettaGridVecEnvPair needs an even number of envs")
        kw.pop("host", None)
        half = num_envs // 2
        self.halves = [GpuMettaGridVecEnv(cfg, half, seed=seed, host=True, first_env=h * half, **kw) for h in range(2)]
        first = self.halves[0]
        self.num_envs, self.agents_per_env = num_envs, first.agents_per_env
        self.num_agents = num_envs * first.agents_per_env
        self.batch_size = self.num_agents // 2  # agents per recv()
        self.single_observation_space, self.single_action_space = first.single_observation_space, first.single_action_space
        self.action_names = first.action_names
        self.driver_env = first
        self._next = self._last = 0

    def async_reset(self, seed: Optional[int] = None) -> None:
        for h in self.halves:
            h.async_reset(seed)
        self._next = self._last = 0

    def recv(self):
        self._last = self._next
        self._next ^= 1
        return self.halves[self._last].recv()

    def send(self, actions) -> None:
        self.halves[self._last].send(actions)

    def close(self) -> None:
        for h in self.halves:
            h.close()

    def episode_stats(self) -> dict[str, float]:
        a, b = (h.episode_stats() for h in self.halves)
        return {k: a.get(k, 0.0) + b.get(k, 0.0) for k in set(a) | set(b)}
