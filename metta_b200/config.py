"""MettaGridConfig schema, host side.

Field-for-field mirror of the reference's pydantic schema so that a config built for the
reference (``MettaGridConfig.model_dump()``) loads here unchanged and vice versa.

Reference: packages/mettagrid/python/src/mettagrid/config/mettagrid_config.py:30-613
(AgentRewards ... MettaGridConfig) and config/obs_config.py:13-25.

Only the schema is mirrored: validation rules that the simulator relies on (tag defaults,
collective tags, vibe_names sync, damage key overlap) are kept; map builders, renderers and
policies are out of scope, so ``map_builder`` is carried as opaque data and maps are handed to
the simulator as already-built cell-name grids (see ``metta_b200.maps``).
"""

from __future__ import annotations

from typing import Any, Literal, Optional

from pydantic import BaseModel, ConfigDict, Field, model_validator

Direction = Literal["north", "south", "east", "west", "northeast", "northwest", "southeast", "southwest"]
CardinalDirections = ["north", "south", "west", "east"]


class Config(BaseModel):
    """Base: unknown keys are errors, like the reference's ``Config`` (base_config.py:9-17)."""

    model_config = ConfigDict(extra="forbid")

    def override(self, key: str, value: Any) -> "Config":
        """Dotted-path override (reference base_config.py:45-139), dict keys may contain dots."""
        parts = key.split(".")
        node: Any = self
        i = 0
        while i < len(parts) - 1:
            part = parts[i]
            if isinstance(node, dict):
                if part in node:
                    node = node[part]
                    i += 1
                    continue
                rest = ".".join(parts[i:])
                node[rest] = value
                return self
            if not hasattr(node, part):
                raise ValueError(f"Override {key} failed: no field {part!r}")
            node = getattr(node, part)
            i += 1
        last = parts[-1]
        if isinstance(node, dict):
            node[last] = value
        else:
            if not hasattr(node, last):
                raise ValueError(f"Override {key} failed: no field {last!r}")
            setattr(node, last, value)
        return self


class Vibe(Config):
    """One entry of the vibe catalogue (reference config/vibes.py:6-13)."""

    symbol: str = ""
    name: str
    category: str = "misc"


# The reference ships a ~150 entry emoji catalogue (config/vibes.py:35-196). It is content, not
# logic: only the *position* of a name in ``change_vibe.vibes`` matters to the simulator. Native
# configs start from the neutral vibe only; reference configs bring their own list.
DEFAULT_VIBES = [Vibe(symbol="", name="default", category="emotion")]


class AgentRewards(Config):
    inventory: dict[str, float] = Field(default_factory=dict)
    inventory_max: dict[str, float] = Field(default_factory=dict)
    stats: dict[str, float] = Field(default_factory=dict)
    stats_max: dict[str, float] = Field(default_factory=dict)


class ResourceLimitsConfig(Config):
    limit: int
    resources: list[str]
    modifiers: dict[str, int] = Field(default_factory=dict)


class InventoryConfig(Config):
    default_limit: int = Field(default=65535, ge=0)
    limits: dict[str, ResourceLimitsConfig] = Field(default_factory=dict)
    initial: dict[str, int] = Field(default_factory=dict)
    regen_amounts: dict[str, dict[str, int]] = Field(default_factory=dict)

    def get_limit(self, resource_name: str) -> int:
        for limit_config in self.limits.values():
            if resource_name in limit_config.resources:
                return limit_config.limit
        return self.default_limit


class DamageConfig(Config):
    threshold: dict[str, int] = Field(default_factory=dict)
    resources: dict[str, int] = Field(default_factory=dict)

    @model_validator(mode="after")
    def _distinct_keys(self) -> "DamageConfig":
        both = set(self.threshold) & set(self.resources)
        if both:
            raise ValueError(
                f"Resources cannot appear in both threshold and resources maps. Overlapping keys: {sorted(both)}"
            )
        return self


class AgentConfig(Config):
    inventory: InventoryConfig = Field(default_factory=InventoryConfig)
    rewards: AgentRewards = Field(default_factory=AgentRewards)
    freeze_duration: int = Field(default=10, ge=-1)
    team_id: int = Field(default=0, ge=0)
    tags: list[str] = Field(default_factory=lambda: ["agent"])
    diversity_tracked_resources: list[str] = Field(default_factory=list)
    initial_vibe: int = Field(default=0, ge=0)
    damage: Optional[DamageConfig] = None


class ActionConfig(Config):
    action_handler: str
    enabled: bool = True
    required_resources: dict[str, int] = Field(default_factory=dict)
    consumed_resources: dict[str, int] = Field(default_factory=dict)

    def action_names(self) -> list[str]:
        return self._action_names() if self.enabled else []

    def _action_names(self) -> list[str]:
        return []


class NoopActionConfig(ActionConfig):
    action_handler: str = "noop"

    def _action_names(self) -> list[str]:
        return ["noop"]


class MoveActionConfig(ActionConfig):
    action_handler: str = "move"
    allowed_directions: list[Direction] = Field(default_factory=lambda: list(CardinalDirections))

    def _action_names(self) -> list[str]:
        return [f"move_{d}" for d in self.allowed_directions]


class ChangeVibeActionConfig(ActionConfig):
    action_handler: str = "change_vibe"
    vibes: list[Vibe] = Field(default_factory=lambda: list(DEFAULT_VIBES))

    def _action_names(self) -> list[str]:
        return [f"change_vibe_{v.name}" for v in self.vibes]


class AttackOutcome(Config):
    actor_inv_delta: dict[str, int] = Field(default_factory=dict)
    target_inv_delta: dict[str, int] = Field(default_factory=dict)
    loot: list[str] = Field(default_factory=list)
    freeze: int = 0


class AttackActionConfig(ActionConfig):
    action_handler: str = "attack"
    defense_resources: dict[str, int] = Field(default_factory=dict)
    armor_resources: dict[str, int] = Field(default_factory=dict)
    weapon_resources: dict[str, int] = Field(default_factory=dict)
    success: AttackOutcome = Field(default_factory=AttackOutcome)
    vibes: list[str] = Field(default_factory=list)
    vibe_bonus: dict[str, int] = Field(default_factory=dict)


class VibeTransfer(Config):
    vibe: str
    target: dict[str, int] = Field(default_factory=dict)
    actor: dict[str, int] = Field(default_factory=dict)


class TransferActionConfig(ActionConfig):
    action_handler: str = "transfer"
    vibe_transfers: list[VibeTransfer] = Field(default_factory=list)


class ActionsConfig(Config):
    noop: NoopActionConfig = Field(default_factory=NoopActionConfig)
    move: MoveActionConfig = Field(default_factory=MoveActionConfig)
    attack: AttackActionConfig = Field(default_factory=lambda: AttackActionConfig(enabled=False))
    transfer: TransferActionConfig = Field(default_factory=lambda: TransferActionConfig(enabled=False))
    change_vibe: ChangeVibeActionConfig = Field(default_factory=ChangeVibeActionConfig)

    def action_names(self) -> list[str]:
        """Action index order: noop, moves, (attack, transfer: none), change_vibe_*.

        Reference: mettagrid_config.py:323-327 and mettagrid_c.cpp:291-352.
        """
        out: list[str] = []
        for a in (self.noop, self.move, self.attack, self.transfer, self.change_vibe):
            out += a.action_names()
        return out


class GlobalObsConfig(Config):
    episode_completion_pct: bool = True
    last_action: bool = True
    last_reward: bool = True
    compass: bool = False
    goal_obs: bool = False


class ObsConfig(Config):
    width: int = 11
    height: int = 11
    token_dim: int = 3
    num_tokens: int = 200
    token_value_base: int = 256


class GridObjectConfig(Config):
    name: str
    map_name: str = ""
    render_name: str = ""
    render_symbol: str = "?"
    tags: list[str] = Field(default_factory=list)
    vibe: int = Field(default=0, ge=0, le=255)
    collective: Optional[str] = None

    @model_validator(mode="after")
    def _defaults_from_name(self) -> "GridObjectConfig":
        if not self.map_name:
            self.map_name = self.name
        if not self.render_name:
            self.render_name = self.name
        if not self.tags:
            self.tags = [self.render_name]
        if self.collective:
            tag = f"collective:{self.collective}"
            if tag not in self.tags:
                self.tags = self.tags + [tag]
        return self


class WallConfig(GridObjectConfig):
    pydantic_type: Literal["wall"] = "wall"
    name: str = "wall"


class ProtocolConfig(Config):
    min_agents: int = Field(default=0, ge=0)
    vibes: list[str] = Field(default_factory=list)
    input_resources: dict[str, int] = Field(default_factory=dict)
    output_resources: dict[str, int] = Field(default_factory=dict)
    cooldown: int = Field(default=0, ge=0)


class AssemblerConfig(GridObjectConfig):
    pydantic_type: Literal["assembler"] = "assembler"
    protocols: list[ProtocolConfig] = Field(default_factory=list)
    allow_partial_usage: bool = False
    max_uses: int = Field(default=0, ge=0)
    clip_immune: bool = False
    start_clipped: bool = False
    chest_search_distance: int = Field(default=0, ge=0)


class ChestConfig(GridObjectConfig):
    pydantic_type: Literal["chest"] = "chest"
    name: str = "chest"
    vibe_transfers: dict[str, dict[str, int]] = Field(default_factory=dict)
    inventory: InventoryConfig = Field(default_factory=InventoryConfig)


class ClipperConfig(Config):
    unclipping_protocols: list[ProtocolConfig] = Field(default_factory=list)
    length_scale: int = Field(default=0, ge=0)
    scaled_cutoff_distance: int = Field(default=3, ge=1)
    clip_period: int = Field(default=0, ge=0)


class CollectiveConfig(Config):
    name: str
    inventory: InventoryConfig = Field(default_factory=InventoryConfig)


_OBJECT_CLASSES = {"wall": WallConfig, "assembler": AssemblerConfig, "chest": ChestConfig}

DEFAULT_RESOURCES = [
    "ore_red",
    "ore_blue",
    "ore_green",
    "battery_red",
    "battery_blue",
    "battery_green",
    "heart",
    "armor",
    "laser",
    "blueprint",
]


class GameConfig(Config):
    resource_names: list[str] = Field(default_factory=lambda: list(DEFAULT_RESOURCES))
    vibe_names: list[str] = Field(default_factory=list)
    num_agents: int = Field(default=24, ge=1)
    max_steps: int = Field(default=10000, ge=0)
    episode_truncates: bool = False
    obs: ObsConfig = Field(default_factory=ObsConfig)
    agent: AgentConfig = Field(default_factory=AgentConfig)
    agents: list[AgentConfig] = Field(default_factory=list)
    actions: ActionsConfig = Field(default_factory=ActionsConfig)
    global_obs: GlobalObsConfig = Field(default_factory=GlobalObsConfig)
    objects: dict[str, Any] = Field(default_factory=dict)
    params: Optional[Any] = None
    inventory_regen_interval: int = Field(default=0, ge=0)
    clipper: Optional[ClipperConfig] = None
    collectives: list[CollectiveConfig] = Field(default_factory=list)
    map_builder: Optional[Any] = None
    protocol_details_obs: bool = True
    reward_estimates: Optional[dict[str, float]] = None

    @model_validator(mode="before")
    @classmethod
    def _strip_computed(cls, data: Any) -> Any:
        # model_dump() of the reference emits the computed ``obs.features`` list; drop it like
        # convert_to_cpp_game_config does (mettagrid_c_config.py:37-44).
        if isinstance(data, dict) and isinstance(data.get("obs"), dict) and "features" in data["obs"]:
            data = dict(data)
            data["obs"] = {k: v for k, v in data["obs"].items() if k != "features"}
        return data

    @model_validator(mode="after")
    def _resolve(self) -> "GameConfig":
        objs: dict[str, GridObjectConfig] = {}
        for key, val in self.objects.items():
            if isinstance(val, GridObjectConfig):
                objs[key] = val
            elif isinstance(val, dict):
                kind = val.get("pydantic_type")
                if kind not in _OBJECT_CLASSES:
                    raise ValueError(f"Unknown object type: {key}")
                objs[key] = _OBJECT_CLASSES[kind](**val)
            elif hasattr(val, "model_dump"):
                d = val.model_dump()
                objs[key] = _OBJECT_CLASSES[d["pydantic_type"]](**d)
            else:
                raise ValueError(f"Unknown object type: {key}")
        self.objects = objs
        self.vibe_names = [v.name for v in self.actions.change_vibe.vibes]
        return self


class MettaGridConfig(Config):
    label: str = "mettagrid"
    game: GameConfig = Field(default_factory=GameConfig)
    desync_episodes: bool = True

    @staticmethod
    def from_reference(cfg: Any) -> "MettaGridConfig":
        """Accept the reference's own MettaGridConfig (or its dump) unchanged."""
        if isinstance(cfg, MettaGridConfig):
            return cfg
        d = cfg if isinstance(cfg, dict) else cfg.model_dump()
        return MettaGridConfig(**d)
