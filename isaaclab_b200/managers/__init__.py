"""Manager term-config API (Contract 1 of SURVEY.md 8b): the cfg classes of LAB/managers/manager_term_cfg.py:28-344
and `SceneEntityCfg` (LAB/managers/scene_entity_cfg.py) with the same field names, so task cfgs written for the
reference read unchanged.  They are plain dataclasses; `isaaclab_b200.envs.compile` lowers them to fused opcodes."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable


@dataclass
class SceneEntityCfg:
    name: str = "robot"
    joint_names: str | list[str] | None = None
    joint_ids: list[int] | slice = field(default_factory=lambda: slice(None))
    body_names: str | list[str] | None = None
    body_ids: list[int] | slice = field(default_factory=lambda: slice(None))
    preserve_order: bool = False

    def resolve(self, scene):
        """LAB/managers/scene_entity_cfg.py:112-254: names -> ids (or slice(None) when every entity is selected)."""
        if self.name not in scene.keys():
            raise ValueError(f"The scene entity '{self.name}' does not exist. Available entities: {list(scene.keys())}.")
        entity = scene[self.name]
        if self.joint_names is not None:
            if not hasattr(entity, "find_joints"):
                raise ValueError(f"The scene entity '{self.name}' has no joints.")
            ids, _ = entity.find_joints(self.joint_names, preserve_order=self.preserve_order)
            self.joint_ids = slice(None) if len(ids) == entity.num_joints and not self.preserve_order else ids
        if self.body_names is not None:
            ids, _ = entity.find_bodies(self.body_names, preserve_order=self.preserve_order)
            self.body_ids = slice(None) if len(ids) == entity.num_bodies and not self.preserve_order else ids


class ManagerTermBase:
    """Base of class-type terms (LAB/managers/manager_base.py:28-115): constructed once with `(cfg, env)` when the env is
    built, called as `term(env, **params)` every step, `reset(env_ids)` for the envs that were reset."""

    def __init__(self, cfg, env):
        self.cfg = cfg
        self._env = env

    @property
    def num_envs(self) -> int:
        return self._env.num_envs

    @property
    def device(self):
        return self._env.device

    def reset(self, env_ids=None) -> None:
        pass

    def __call__(self, *args):
        raise NotImplementedError("The method '__call__' should be implemented by the term.")


@dataclass
class ManagerTermBaseCfg:
    func: Callable | str = None
    params: dict[str, Any] = field(default_factory=dict)


@dataclass
class NoiseCfg:
    operation: str = "add"


@dataclass
class UniformNoiseCfg(NoiseCfg):
    n_min: float = -1.0
    n_max: float = 1.0


@dataclass
class ConstantNoiseCfg(NoiseCfg):
    bias: float = 0.0


@dataclass
class GaussianNoiseCfg(NoiseCfg):
    mean: float = 0.0
    std: float = 1.0


@dataclass
class ObservationTermCfg(ManagerTermBaseCfg):
    modifiers: list | None = None
    noise: NoiseCfg | None = None
    clip: tuple[float, float] | None = None
    scale: float | None = None
    history_length: int = 0
    flatten_history_dim: bool = True


@dataclass
class ObservationGroupCfg:
    concatenate_terms: bool = True
    enable_corruption: bool = False
    history_length: int | None = None
    flatten_history_dim: bool = True


@dataclass
class RewardTermCfg(ManagerTermBaseCfg):
    weight: float = None


@dataclass
class TerminationTermCfg(ManagerTermBaseCfg):
    time_out: bool = False


@dataclass
class EventTermCfg(ManagerTermBaseCfg):
    mode: str = None
    interval_range_s: tuple[float, float] | None = None
    is_global_time: bool = False
    min_step_count_between_reset: int = 0


@dataclass
class CurriculumTermCfg(ManagerTermBaseCfg):
    pass


@dataclass
class ActionTermCfg:
    class_type: Any = None
    asset_name: str = "robot"
    debug_vis: bool = False
    clip: dict | None = None


@dataclass
class CommandTermCfg:
    class_type: Any = None
    resampling_time_range: tuple[float, float] = (10.0, 10.0)
    debug_vis: bool = False


def term_items(cfg_obj):
    """(name, term_cfg) pairs of a cfg container in declaration order, skipping None entries (how the reference's
    managers iterate `cfg.__dict__.items()`, LAB/managers/reward_manager.py:214-236)."""
    out = []
    for klass in reversed(type(cfg_obj).__mro__):
        for name in vars(klass):
            if name.startswith("_") or name in [n for n, _ in out]:
                continue
            out.append((name, None))
    for name in vars(cfg_obj):
        if not name.startswith("_") and name not in [n for n, _ in out]:
            out.append((name, None))
    res = []
    for name, _ in out:
        v = getattr(cfg_obj, name, None)
        if v is None or callable(v) and not isinstance(v, (ManagerTermBaseCfg, ActionTermCfg, CommandTermCfg)):
            continue
        if isinstance(v, (ManagerTermBaseCfg, ActionTermCfg, CommandTermCfg)):
            res.append((name, v))
    return res
