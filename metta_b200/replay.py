"""Replay logs of one env of a batch, in the reference's replay format (version 3): the slow path beside a
rollout that mettascope and the reference's tooling read.

Reference: `mettagrid/simulator/replay_log_writer.py` (EpisodeReplay :96-226, ReplayLogWriter :48-93) and
`mettagrid/util/grid_object_formatter.py` (what is kept of every grid object). The file is
zlib-compressed JSON: a header (names, map size, the config) and one record per grid object whose
entries are change lists `[[step, value], ...]` -- a key is only extended when its value changes, a key
first seen after step 0 starts with `[0, 0]`, a key that vanishes gets a `[step, 0]`, and a key that never
changed is collapsed to its value when the replay is finished.

Nothing here is on the step path: every logged step reads the env's records back from the device
(`grid_objects`), as the reference's writer calls `MettaGrid::grid_objects()` every step.

    writer = ReplayLogWriter(vecenv, "/tmp/replays", env_index=0)
    obs, _ = vecenv.reset()
    while ...:
        obs, rew, term, trunc, infos = vecenv.step(actions)
        writer.on_step(actions)          # writes <dir>/<uuid>.json.z when the env's episode ends
"""

from __future__ import annotations

import json
import os
import uuid
import zlib
from typing import Any, Optional

import numpy as np

from .batch import Batch
from .grid_objects import grid_objects

# objects of these types never change: recorded at step 0 only (replay_log_writer.py:99-100)
STATIC_TYPES = ("wall",)


def _pairs(d: Optional[dict]) -> list:
    return [] if not d else [[k, v] for k, v in d.items()]


def replay_record(obj: dict, actions, success, rewards, totals) -> dict:
    """What the replay keeps of one grid_objects() entry (grid_object_formatter.py:8-155)."""
    rec: dict[str, Any] = {
        "id": obj["id"],
        "type_name": obj["type_name"],
        "location": list(obj["location"]),
        "orientation": obj.get("orientation", 0),
        "inventory": _pairs(obj.get("inventory")),
        "inventory_max": obj.get("inventory_max", 0),
        "color": obj.get("color", 0),
    }
    if "collective_id" in obj:
        rec["collective_id"] = obj["collective_id"]
    if "agent_id" in obj:
        k = obj["agent_id"]
        rec["is_agent"] = True
        rec["agent_id"] = k
        rec["vision_size"] = 11
        rec["action_id"] = int(np.asarray(actions[k]).reshape(-1)[0])  # flat action index; no parameter
        rec["action_param"] = 0
        rec["action_success"] = bool(success[k])
        rec["current_reward"] = rewards[k].item()
        rec["total_reward"] = totals[k].item()
        rec["freeze_remaining"] = obj.get("freeze_remaining", 0)
        rec["is_frozen"] = obj.get("is_frozen", False)
        rec["freeze_duration"] = obj.get("freeze_duration", 0)
        rec["group_id"] = obj["group_id"]
        rec["vibe_id"] = obj.get("vibe", 0)
        rec["vibe"] = obj.get("vibe", 0)
    elif "protocols" in obj:
        for key, default in (("cooldown_remaining", 0), ("cooldown_duration", 0), ("is_clipped", False), ("is_clip_immune", False),
                             ("uses_count", 0), ("max_uses", 0), ("allow_partial_usage", False)):
            rec[key] = obj.get(key, default)
        rec["protocols"] = [
            {"minAgents": p.get("min_agents", 0), "vibes": list(p.get("vibes", [])), "inputs": _pairs(p.get("inputs")),
             "outputs": _pairs(p.get("outputs")), "cooldown": p.get("cooldown", 0)}
            for p in obj.get("protocols", [])
        ]
    return rec


class EpisodeReplay:
    """The replay of one episode of env `env_index` of a batch."""

    def __init__(self, batch: Batch, env_index: int, mg_config: Optional[dict] = None):
        low = batch.lowered
        self.batch, self.env = batch, env_index
        self.step = 0
        self.objects: list[dict] = []
        self._slot: dict[int, int] = {}  # object id -> index in self.objects
        self.total_rewards = np.zeros(batch.num_agents)
        names = low.type_names
        # one entry per type id, and at least one per object config the core is given: the map cell names
        # (mettagrid_c.cpp:163-178 sizes the list by game_config.objects first)
        n_types = max(len(low.cell_codes), max(names) + 1) if names else 0
        H, W = batch.pool.codes.shape[1:]
        self.header: dict[str, Any] = {
            "version": 3,
            "action_names": list(low.action_names),
            "item_names": list(low.resource_names),
            "type_names": [names.get(i, "") for i in range(n_types)],
            "map_size": [int(W), int(H)],
            "num_agents": int(batch.num_agents),
            "max_steps": 0,  # becomes the number of steps logged (data())
            "mg_config": mg_config,
            "objects": self.objects,
        }

    def log_step(self, current_step: int, actions, rewards, success) -> None:
        """After a step of the env: `actions` i32[A] as sent, `rewards` f32[A] and `success` bool[A] as returned."""
        rewards = np.asarray(rewards)
        self.total_rewards += rewards
        objs = grid_objects(self.batch, self.env, ignore_types=() if self.step == 0 else STATIC_TYPES)
        for oid, obj in objs.items():
            k = self._slot.get(oid)
            if k is None:
                k = self._slot[oid] = len(self.objects)
                self.objects.append({})
            self._merge(self.objects[k], replay_record(obj, actions, success, rewards, self.total_rewards))
        self.step += 1
        if current_step != self.step:
            raise ValueError(f"replay of env {self.env}: step {current_step} logged as step {self.step} (a step was skipped or logged twice)")

    def _merge(self, series: dict, now: dict) -> None:
        t = self.step
        for key, value in now.items():
            s = series.get(key)
            if s is None:
                series[key] = [[t, value]] if t == 0 else [[0, 0], [t, value]]
            elif s[-1][1] != value:
                s.append([t, value])
        for key, s in series.items():
            if key not in now and s[-1][1] != 0:
                s.append([t, 0])

    def data(self) -> dict:
        """The finished replay as plain Python containers (keys that never changed collapse to their value)."""
        self.header["max_steps"] = self.step
        for series in self.objects:
            for key, s in list(series.items()):
                if isinstance(s, list) and len(s) == 1 and isinstance(s[0], (list, tuple)) and len(s[0]) == 2:
                    series[key] = s[0][1]
        return self.header

    def compressed(self) -> bytes:
        return zlib.compress(json.dumps(self.data()).encode("utf-8"))

    def write(self, path: str) -> None:
        with open(path, "wb") as f:
            f.write(self.compressed())


class ReplayLogWriter:
    """Records env `env_index` of a GpuMettaGridVecEnv episode after episode: call on_step(actions) after every
    vecenv.step(actions); a file `<replay_dir>/<uuid>.json.z` is written whenever the env's episode ends."""

    def __init__(self, vecenv, replay_dir: str, env_index: int = 0):
        self.vecenv, self.dir, self.env = vecenv, replay_dir, env_index
        self.paths: list[str] = []
        self._cfg = vecenv.cfg.model_dump(mode="json") if hasattr(vecenv, "cfg") else None
        self._new_episode()

    def _new_episode(self) -> None:
        self.episode_id = str(uuid.uuid4())
        self.replay = EpisodeReplay(self.vecenv.batch, self.env, self._cfg)

    def on_step(self, actions) -> None:
        v, A, i = self.vecenv, self.vecenv.batch.num_agents, self.env
        if getattr(v, "_inflight", False):
            v._wait()  # host mode: the step that send() enqueued
        to_np = lambda x: x.detach().cpu().numpy() if hasattr(x, "detach") else np.asarray(x)  # noqa: E731
        sl = slice(i * A, (i + 1) * A)
        rewards = to_np(v._h_rew if v.host else v.rewards)[sl]
        term = to_np(v._h_term if v.host else v.terminals)[sl]
        trunc = to_np(v._h_trunc if v.host else v.truncations)[sl]
        # the env's own step count (host mirror): a step that was not logged shows up as a mismatch, as in the reference
        self.replay.log_step(int(v._steps[i]), to_np(actions).reshape(-1)[sl], rewards, v.action_success()[i])
        if term.all() or trunc.all():
            path = os.path.join(self.dir, self.episode_id + ".json.z")
            self.replay.write(path)
            self.paths.append(path)
            self._new_episode()
