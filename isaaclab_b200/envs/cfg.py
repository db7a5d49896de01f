"""Env / scene / sensor cfg containers with the reference's field names (LAB/envs/manager_based_rl_env_cfg.py,
LAB/sensors/ray_caster/ray_caster_cfg.py, patterns_cfg.py, LAB/sensors/contact_sensor/contact_sensor_cfg.py,
LAB/terrains/terrain_importer_cfg.py) -- the subset the non-physics step reads."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any

from ..assets import ArticulationCfg


@dataclass
class GridPatternCfg:
    resolution: float = None
    size: tuple[float, float] = None
    direction: tuple[float, float, float] = (0.0, 0.0, -1.0)
    ordering: str = "xy"


@dataclass
class RayCasterCfg:
    @dataclass
    class OffsetCfg:
        pos: tuple = (0.0, 0.0, 0.0)
        rot: tuple = (1.0, 0.0, 0.0, 0.0)

    prim_path: str = ""
    update_period: float = 0.0
    offset: "RayCasterCfg.OffsetCfg" = field(default_factory=lambda: RayCasterCfg.OffsetCfg())
    attach_yaw_only: bool = False
    pattern_cfg: Any = None
    max_distance: float = 1e6
    drift_range: tuple[float, float] = (0.0, 0.0)
    mesh_prim_paths: list = field(default_factory=lambda: ["/World/ground"])
    debug_vis: bool = False


@dataclass
class ContactSensorCfg:
    prim_path: str = ""
    update_period: float = 0.0
    history_length: int = 0
    track_air_time: bool = False
    force_threshold: float = 1.0
    debug_vis: bool = False


@dataclass
class TerrainImporterCfg:
    terrain_type: str = "generator"       # "generator" | "plane"
    terrain_generator: Any = None         # isaaclab_b200.terrains.TerrainGeneratorCfg
    max_init_terrain_level: int | None = 5
    env_spacing: float | None = None


@dataclass
class SimCfg:
    dt: float = 1.0 / 60.0
    render_interval: int = 1


@dataclass
class SceneCfg:
    num_envs: int = 4096
    env_spacing: float = 2.5
    robot: ArticulationCfg = None
    terrain: TerrainImporterCfg | None = None
    height_scanner: RayCasterCfg | None = None
    contact_forces: ContactSensorCfg | None = None


@dataclass
class ManagerBasedRLEnvCfg:
    scene: SceneCfg = None
    observations: Any = None
    actions: Any = None
    commands: Any = None
    rewards: Any = None
    terminations: Any = None
    events: Any = None
    curriculum: Any = None
    decimation: int = 1
    episode_length_s: float = 10.0
    sim: SimCfg = field(default_factory=SimCfg)
    seed: int | None = None
    is_finite_horizon: bool = False
