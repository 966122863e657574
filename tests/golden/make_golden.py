"""Generates tests/golden/*.npz + *.json by RUNNING THE REFERENCE in the build container.

Needs /root/reference (Python package) and oracle/_ref (compiled core). Run:
    python tests/golden/make_golden.py
Each fixture holds the config (reference schema, JSON), the map (cell names), the env seed, the
action sequence and everything the reference produced: observations (initial + every step),
rewards, terminals, truncations, action_success, episode rewards and the final inventories in
the reference's own iteration order (grid_objects()[id]["inventory"]).
The arena fixture is produced through the reference's OWN Python lowering
(mettagrid_c_config.convert_to_cpp_game_config) and map builder; the others through
oracle/ref_bridge.py, which tests/test_oracle_vs_reference.py checks against that lowering.
"""

import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.ref_loader import load_reference_python  # noqa: E402

load_reference_python()

from harness import random_actions  # noqa: E402
from scenarios import SCENARIOS  # noqa: E402

from metta_b200.config import MettaGridConfig  # noqa: E402
from metta_b200.lowering import lower_game_config  # noqa: E402
from oracle.ref_bridge import RefEnv  # noqa: E402


def run(name, cfg, grid, seed, steps, c_cfg=None, restarts=False, invalid_every=0):
    lowered = lower_game_config(cfg.game, grid.shape[0], grid.shape[1])
    env = RefEnv(cfg.game, grid.tolist(), seed, c_cfg=c_cfg)
    A = cfg.game.num_agents
    rng = np.random.RandomState(seed + 1000)
    obs = [env.observations.copy()]
    rew, term, trunc, succ, acts, eprew = [], [], [], [], [], []
    for s in range(steps):
        if restarts and (env.terminals.all() or env.truncations.all()):
            env = RefEnv(cfg.game, grid.tolist(), seed, c_cfg=c_cfg)  # mettagrid_puffer_env.py:158-164
        a = random_actions(rng, cfg, lowered, A, invalid_every=invalid_every, step=s)
        env.step(a)
        acts.append(a)
        obs.append(env.observations.copy())
        rew.append(env.rewards.copy())
        term.append(env.terminals.copy())
        trunc.append(env.truncations.copy())
        succ.append(env.action_success())
        eprew.append(env.episode_rewards().copy())
    objs = env.grid_objects()
    inv_items = np.full((A, 16), -1, np.int32)
    inv_amounts = np.zeros((A, 16), np.int32)
    pos = np.zeros((A, 2), np.int32)
    for o in objs.values():
        if "agent_id" in o:
            k = o["agent_id"]
            pos[k] = (o["r"], o["c"])
            for j, (item, amount) in enumerate(o["inventory"].items()):
                inv_items[k, j] = item
                inv_amounts[k, j] = amount
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"),
        grid=grid, seed=np.int64(seed), actions=np.array(acts, np.int32), obs=np.array(obs, np.uint8),
        rewards=np.array(rew, np.float32), terminals=np.array(term), truncations=np.array(trunc),
        success=np.array(succ), episode_rewards=np.array(eprew, np.float32),
        final_pos=pos, final_inv_items=inv_items, final_inv_amounts=inv_amounts, restarts=np.bool_(restarts),
    )
    with open(os.path.join(HERE, name + ".json"), "w") as f:
        json.dump(cfg.model_dump(), f, indent=1, sort_keys=False, default=str)
    print(name, "steps", steps, "obs", np.array(obs).shape, "nonzero rewards", int((np.array(rew) != 0).sum()))


def main():
    # arena: the reference's own config, lowering and map builder
    from mettagrid.builder.envs import make_arena
    from mettagrid.config.mettagrid_c_config import convert_to_cpp_game_config

    ref_cfg = make_arena(24)
    ref_cfg.game.map_builder.seed = 1000
    grid = ref_cfg.game.map_builder.create().build_for_num_agents(24).grid
    c_cfg = convert_to_cpp_game_config(ref_cfg.game.model_dump())
    dump = ref_cfg.model_dump()
    dump["game"]["map_builder"] = None  # map builders are out of scope; the grid is stored instead
    cfg = MettaGridConfig.from_reference(dump)
    run("arena", cfg, grid, 42, 120, c_cfg=c_cfg)

    for name, steps, restarts, inv in (("kitchen_sink", 250, False, 40), ("tiny_truncating", 23, True, 0), ("clipped_start", 150, False, 0)):
        cfg, grid = SCENARIOS[name]()
        run(name, cfg, grid, 7, steps, restarts=restarts, invalid_every=inv)


def cogames():
    """cogames missions (BASELINE configs[3]): the mission's own MettaGridConfig (157 actions: noop, 4
    moves, 152 vibes; assemblers keyed by vibes, chests, chargers, regen, clipper with clip_period 0)
    through the reference's own lowering and its own map builder."""
    sys.path.insert(0, "/root/reference/packages/cogames/src")
    from cogames.cogs_vs_clips import missions as M
    from mettagrid.config.mettagrid_c_config import convert_to_cpp_game_config

    for name, mission, cogs, steps in (
        ("cogames_vibe_check", M.VibeCheckMission, 4, 300),
        ("cogames_machina_chests", M.Machina1OpenWorldWithChestsMission, 4, 200),
    ):
        ref_cfg = mission.model_copy(update={"num_cogs": cogs}).make_env()
        grid = ref_cfg.game.map_builder.create().build_for_num_agents(ref_cfg.game.num_agents).grid
        c_cfg = convert_to_cpp_game_config(ref_cfg.game.model_dump())
        dump = ref_cfg.model_dump()
        dump["game"]["map_builder"] = None
        cfg = MettaGridConfig.from_reference(dump)
        run(name, cfg, grid, 5, steps, c_cfg=c_cfg)


def cogames_clipper_config():
    """RepairMission (clipper with clip_period 50): config and map only. Which assembler the reference's
    clipper picks depends on heap addresses, so its outputs are not a fixture; the config
    is used for oracle <-> kernel lockstep."""
    sys.path.insert(0, "/root/reference/packages/cogames/src")
    from cogames.cogs_vs_clips import missions as M

    ref_cfg = M.RepairMission.model_copy(update={"num_cogs": 4}).make_env()
    grid = ref_cfg.game.map_builder.create().build_for_num_agents(ref_cfg.game.num_agents).grid
    dump = ref_cfg.model_dump()
    dump["game"]["map_builder"] = None
    cfg = MettaGridConfig.from_reference(dump)
    np.savez_compressed(os.path.join(HERE, "cogames_repair_config.npz"), grid=grid)
    with open(os.path.join(HERE, "cogames_repair_config.json"), "w") as f:
        json.dump(cfg.model_dump(), f, indent=1, sort_keys=False, default=str)
    print("cogames_repair_config", grid.shape, "clip_period", cfg.game.clipper.clip_period if cfg.game.clipper else None)


def arena_mapgen_pool(n=256):
    """BASELINE configs[1] asks for arena maps from the reference's MapGen (builder/envs.py:57-75,
    map_builder.seed = 1000 + env index). The builders are out of scope; their output is
    committed: `names` (the cell names that occur) and `cells` u8 [n][62][62] indexing into it."""
    from mettagrid.builder.envs import make_arena

    grids = []
    for i in range(n):
        ref_cfg = make_arena(24)
        ref_cfg.game.map_builder.seed = 1000 + i
        grids.append(ref_cfg.game.map_builder.create().build_for_num_agents(24).grid)
    names = sorted({str(x) for g in grids for x in np.unique(g)})
    cells = np.zeros((n,) + grids[0].shape, np.uint8)
    for i, g in enumerate(grids):
        for k, name in enumerate(names):
            cells[i][g == name] = k
    np.savez_compressed(os.path.join(HERE, "arena_mapgen_pool.npz"), names=np.array(names), cells=cells)
    occ = {name: float((cells == k).mean()) for k, name in enumerate(names)}
    print("arena_mapgen_pool", cells.shape, {k: round(v, 4) for k, v in occ.items()})


def clipper():
    """Clipper picks (systems/clipper.hpp:140-238). Must run against the reference build whose clipper sets are
    ordered by object id instead of heap address (oracle/build_ref_ordered.sh, MG_REF_DIR=_ref_ordered): the
    stock build's pick order differs from process to process. Two fixtures: the clipper_spread scenario (about
    400 clip and unclip events in 1500 steps) and a cogames mission with clip_period 50 on its own 110x110 map."""
    assert os.environ.get("MG_REF_DIR") == "_ref_ordered", "run with MG_REF_DIR=_ref_ordered (see oracle/build_ref_ordered.sh)"
    cfg, grid = SCENARIOS["clipper_spread"]()
    run("clipper_spread", cfg, grid, 20, 1500)
    sys.path.insert(0, "/root/reference/packages/cogames/src")
    from cogames.cogs_vs_clips import missions as M
    from mettagrid.config.mettagrid_c_config import convert_to_cpp_game_config

    ref_cfg = M.HelloWorldOpenWorldMission.model_copy(update={"num_cogs": 8, "clip_period": 50}).make_env()
    grid = ref_cfg.game.map_builder.create().build_for_num_agents(ref_cfg.game.num_agents).grid
    c_cfg = convert_to_cpp_game_config(ref_cfg.game.model_dump())
    dump = ref_cfg.model_dump()
    dump["game"]["map_builder"] = None
    cfg = MettaGridConfig.from_reference(dump)
    run("cogames_clip_period", cfg, grid, 3, 700, c_cfg=c_cfg)


def replay():
    """An arena episode of 60 steps recorded by the reference's own replay writer (EpisodeReplay through the
    Simulation event handlers, mettagrid/simulator/replay_log_writer.py): tests/golden/arena_replay.*"""
    import zlib

    import mettagrid.config.mettagrid_config  # noqa: F401  (first: mettagrid.simulator imports it back)
    from mettagrid.builder.envs import make_arena
    from mettagrid.simulator.replay_log_writer import InMemoryReplayWriter
    from mettagrid.simulator.simulator import Simulation

    ref_cfg = make_arena(24)
    ref_cfg.game.map_builder.seed = 1000
    ref_cfg.game.max_steps = 60
    grid = ref_cfg.game.map_builder.create().build_for_num_agents(24).grid
    writer = InMemoryReplayWriter()
    sim = Simulation(ref_cfg, seed=42, event_handlers=[writer])
    rng = np.random.RandomState(77)
    acts = []
    n_actions = len(sim.action_names)
    while not sim.is_done():
        a = rng.randint(0, min(n_actions, 6), size=24).astype(np.int32)
        sim._c_sim.actions()[:] = a
        sim.step()
        acts.append(a)
    (done,) = writer.get_completed_replays()
    data = done.get_replay_data()
    dump = ref_cfg.model_dump()
    dump["game"]["map_builder"] = None
    cfg = MettaGridConfig.from_reference(dump)
    with open(os.path.join(HERE, "arena_replay.json"), "w") as f:
        json.dump(cfg.model_dump(), f, indent=1, sort_keys=False, default=str)
    np.savez_compressed(os.path.join(HERE, "arena_replay.npz"), grid=grid, seed=np.int64(42), actions=np.array(acts, np.int32))
    with open(os.path.join(HERE, "arena_replay.replay.json.z"), "wb") as f:
        f.write(zlib.compress(json.dumps(data).encode("utf-8"), 9))
    print("arena_replay: steps", len(acts), "objects", len(data["objects"]), "bytes", os.path.getsize(os.path.join(HERE, "arena_replay.replay.json.z")))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "replay":
        replay()
    elif len(sys.argv) > 1 and sys.argv[1] == "clipper":
        cogames_clipper_config()
    elif len(sys.argv) > 1 and sys.argv[1] == "cogames":
        cogames()
    elif len(sys.argv) > 1 and sys.argv[1] == "clipper_picks":
        clipper()
    elif len(sys.argv) > 1 and sys.argv[1] == "arena_pool":
        arena_mapgen_pool()
    else:
        main()
        cogames()
        cogames_clipper_config()
        arena_mapgen_pool()

