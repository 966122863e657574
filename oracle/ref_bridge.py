"""metta_b200 GameConfig -> the compiled reference core (oracle/_ref). TEST INFRASTRUCTURE.

Builds the reference's pybind config objects straight from our schema mirror, so the unmodified
reference ``MettaGrid`` can be stepped beside the CUDA simulator wherever the .so travelled,
including the GPU box where /root/reference does not exist. It plays the role of
mettagrid_c_config.py:30-577 for the checker; tests/test_oracle_vs_reference.py verifies (in the
container) that it yields the same observations as the reference's own Python lowering.
"""

from __future__ import annotations

import numpy as np

from metta_b200.config import AssemblerConfig, ChestConfig, GameConfig, WallConfig
from metta_b200.lowering import TEAM_NAMES, compute_features

from .ref_loader import load_mettagrid_c


def _limit_defs(C, inv, rid, resource_names, with_defaults):
    defs, configured = [], set()
    for rl in inv.limits.values():
        ids = [rid[n] for n in rl.resources] if with_defaults else [rid[n] for n in rl.resources if n in rid]
        if not ids:
            continue
        mods = {rid[n]: b for n, b in rl.modifiers.items() if n in rid}
        defs.append(C.LimitDef(ids, rl.limit, mods))
        configured.update(rl.resources)
    if with_defaults:
        for n in resource_names:
            if n not in configured:
                defs.append(C.LimitDef([rid[n]], inv.default_limit))
    ic = C.InventoryConfig()
    ic.limit_defs = defs
    return ic


def _protocol(C, p, rid, vid):
    cp = C.Protocol()
    cp.min_agents = p.min_agents
    cp.vibes = sorted(vid[v] for v in p.vibes)
    cp.input_resources = {int(rid[k]): int(v) for k, v in p.input_resources.items()}
    cp.output_resources = {int(rid[k]): int(v) for k, v in p.output_resources.items()}
    cp.cooldown = p.cooldown
    return cp


def to_reference_config(game: GameConfig):
    C = load_mettagrid_c()
    resource_names = list(game.resource_names)
    rid = {n: i for i, n in enumerate(resource_names)}
    vibes = game.actions.change_vibe.vibes
    vid = {v.name: i for i, v in enumerate(vibes)}
    type_id = {n: i + 1 for i, n in enumerate(sorted(game.objects.keys()))}
    agents = list(game.agents) or [game.agent.model_copy(update={"team_id": 0}) for _ in range(game.num_agents)]
    tags = set()
    for oc in game.objects.values():
        tags.update(oc.tags)
    for ac in agents:
        tags.update(ac.tags)
    tid = {t: i for i, t in enumerate(sorted(tags))}

    objects = {}
    teams: dict[int, list] = {}
    for ac in agents:
        teams.setdefault(ac.team_id, []).append(ac)
    for team_id, members in teams.items():
        first = members[0]
        stat_rewards = dict(first.rewards.stats)
        stat_max = dict(first.rewards.stats_max)
        for k, v in first.rewards.inventory.items():
            stat_rewards[k + ".amount"] = v
        for k, v in first.rewards.inventory_max.items():
            stat_max[k + ".amount"] = v
        dmg = C.DamageConfig()
        if first.damage is not None:
            dmg.threshold = {rid[k]: v for k, v in first.damage.threshold.items()}
            dmg.resources = {rid[k]: v for k, v in first.damage.resources.items()}
        regen = {
            vid[vn]: {rid[k]: v for k, v in amounts.items()} for vn, amounts in first.inventory.regen_amounts.items()
        }
        gname = TEAM_NAMES.get(team_id, f"team_{team_id}")
        ca = C.AgentConfig(
            type_id=0,
            type_name="agent",
            group_id=team_id,
            group_name=gname,
            freeze_duration=first.freeze_duration,
            initial_vibe=first.initial_vibe,
            inventory_config=_limit_defs(C, first.inventory, rid, resource_names, True),
            stat_rewards=stat_rewards,
            stat_reward_max=stat_max,
            initial_inventory={rid[k]: v for k, v in first.inventory.initial.items()},
            inventory_regen_amounts=regen,
            diversity_tracked_resources=[rid[n] for n in first.diversity_tracked_resources if n in rid],
            damage_config=dmg,
        )
        ca.tag_ids = [tid[t] for t in first.tags]
        objects["agent." + gname] = ca
        objects[f"agent.team_{team_id}"] = ca
        if team_id == 0:
            objects["agent.default"] = ca
            objects["agent.agent"] = ca

    for key, oc in game.objects.items():
        name = oc.map_name or key
        if isinstance(oc, WallConfig):
            co = C.WallConfig(type_id=type_id[key], type_name=key, initial_vibe=oc.vibe)
        elif isinstance(oc, AssemblerConfig):
            co = C.AssemblerConfig(type_id=type_id[key], type_name=key, initial_vibe=oc.vibe)
            co.protocols = [_protocol(C, p, rid, vid) for p in reversed(oc.protocols)]
            co.allow_partial_usage = oc.allow_partial_usage
            co.max_uses = oc.max_uses
            co.clip_immune = oc.clip_immune
            co.start_clipped = oc.start_clipped
            co.chest_search_distance = oc.chest_search_distance
        elif isinstance(oc, ChestConfig):
            co = C.ChestConfig(type_id=type_id[key], type_name=key, initial_vibe=oc.vibe)
            co.vibe_transfers = {
                vid[vn]: {rid[r]: d for r, d in deltas.items()} for vn, deltas in oc.vibe_transfers.items()
            }
            co.initial_inventory = {rid[r]: a for r, a in oc.inventory.initial.items()}
            co.inventory_config = _limit_defs(C, oc.inventory, rid, resource_names, False)
        else:
            raise ValueError(key)
        co.tag_ids = [tid[t] for t in oc.tags]
        objects[name] = co

    def action_res(ac):
        if not ac.enabled:
            return {}, {}
        consumed = {rid[k]: int(v) for k, v in ac.consumed_resources.items()}
        src = ac.required_resources or ac.consumed_resources
        return {rid[k]: int(v) for k, v in src.items()}, consumed

    acts = game.actions
    actions = {}
    req, con = action_res(acts.noop)
    actions["noop"] = C.ActionConfig(required_resources=req, consumed_resources=con)
    req, con = action_res(acts.move)
    actions["move"] = C.MoveActionConfig(
        allowed_directions=list(acts.move.allowed_directions), required_resources=req, consumed_resources=con
    )
    req, con = action_res(acts.attack)
    atk = acts.attack
    actions["attack"] = C.AttackActionConfig(
        required_resources=req,
        consumed_resources=con,
        defense_resources={rid[k]: v for k, v in atk.defense_resources.items()},
        armor_resources={rid[k]: v for k, v in atk.armor_resources.items()},
        weapon_resources={rid[k]: v for k, v in atk.weapon_resources.items()},
        success=C.AttackOutcome(
            {rid[k]: v for k, v in atk.success.actor_inv_delta.items()},
            {rid[k]: v for k, v in atk.success.target_inv_delta.items()},
            [rid[n] for n in atk.success.loot],
            atk.success.freeze,
        ),
        enabled=atk.enabled,
        vibes=[vid[v] for v in atk.vibes],
        vibe_bonus={vid[v]: b for v, b in atk.vibe_bonus.items()},
    )
    xf = acts.transfer
    actions["transfer"] = C.TransferActionConfig(
        required_resources={rid[k]: int(v) for k, v in xf.required_resources.items()},
        vibe_transfers={
            vid[vt.vibe]: C.VibeTransferEffect(
                {rid[k]: v for k, v in vt.target.items()}, {rid[k]: v for k, v in vt.actor.items()}
            )
            for vt in xf.vibe_transfers
        },
        enabled=xf.enabled,
    )
    req, con = action_res(acts.change_vibe)
    actions["change_vibe"] = C.ChangeVibeActionConfig(
        required_resources=req,
        consumed_resources=con,
        number_of_vibes=len(vibes) if acts.change_vibe.enabled else 0,
    )

    g = game.global_obs
    kwargs = dict(
        num_agents=game.num_agents,
        max_steps=game.max_steps,
        episode_truncates=game.episode_truncates,
        obs_width=game.obs.width,
        obs_height=game.obs.height,
        resource_names=resource_names,
        vibe_names=[v.name for v in vibes],
        num_observation_tokens=game.obs.num_tokens,
        global_obs=C.GlobalObsConfig(
            episode_completion_pct=g.episode_completion_pct,
            last_action=g.last_action,
            last_reward=g.last_reward,
            compass=g.compass,
            goal_obs=g.goal_obs,
        ),
        feature_ids={name: i for i, name, _ in compute_features(game)},
        actions=actions,
        objects=objects,
        tag_id_map={i: t for t, i in tid.items()},
        protocol_details_obs=game.protocol_details_obs,
        inventory_regen_interval=game.inventory_regen_interval,
        token_value_base=game.obs.token_value_base,
    )
    if game.reward_estimates is not None:
        kwargs["reward_estimates"] = game.reward_estimates
    if game.collectives:  # mettagrid_c_config.py:543-575
        colls = {}
        for c in game.collectives:
            cc_ = C.CollectiveConfig(c.name)
            cc_.inventory_config = _limit_defs(C, c.inventory, rid, resource_names, False)
            cc_.initial_inventory = {rid[k]: v for k, v in c.inventory.initial.items() if k in rid}
            colls[c.name] = cc_
        kwargs["collectives"] = colls
    if game.clipper is not None:
        cc = C.ClipperConfig()
        cc.unclipping_protocols = [_protocol(C, p, rid, vid) for p in game.clipper.unclipping_protocols]
        cc.length_scale = game.clipper.length_scale
        cc.scaled_cutoff_distance = game.clipper.scaled_cutoff_distance
        cc.clip_period = game.clipper.clip_period
        kwargs["clipper"] = cc
    return C.GameConfig(**kwargs)


class RefEnv:
    """The unmodified reference MettaGrid on caller-owned numpy buffers (mettagrid_c.cpp:721-752)."""

    def __init__(self, game: GameConfig, map_names, seed: int, c_cfg=None):
        C = load_mettagrid_c()
        self.c_cfg = c_cfg if c_cfg is not None else to_reference_config(game)
        grid = [list(map(str, row)) for row in map_names]
        self.sim = C.MettaGrid(self.c_cfg, grid, int(seed))
        A, T = game.num_agents, game.obs.num_tokens
        self.observations = np.zeros((A, T, 3), dtype=np.uint8)
        self.terminals = np.zeros(A, dtype=np.bool_)
        self.truncations = np.zeros(A, dtype=np.bool_)
        self.rewards = np.zeros(A, dtype=np.float32)
        self.actions = np.zeros(A, dtype=np.int32)
        self.sim.set_buffers(self.observations, self.terminals, self.truncations, self.rewards, self.actions)

    def step(self, actions=None):
        if actions is not None:
            self.actions[:] = actions
        self.sim.step()

    def episode_rewards(self):
        return np.asarray(self.sim.get_episode_rewards())

    def action_success(self):
        return np.asarray(self.sim.action_success(), dtype=bool)

    def grid_objects(self):
        return self.sim.grid_objects()

    def episode_stats(self):
        return self.sim.get_episode_stats()
