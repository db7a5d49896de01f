"""Locomotion velocity-tracking task cfgs, as data: TASKS/manager_based/locomotion/velocity/velocity_env_cfg.py:66-310
and config/{anymal_c,go1,h1}/rough_env_cfg.py + flat variants, written against isaaclab_b200's cfg classes (same
field names and term identities as the reference)."""
from __future__ import annotations

import copy
import math
from dataclasses import dataclass, field

from ..assets import (ANYMAL_B_CFG, ANYMAL_C_CFG, ANYMAL_D_CFG, CASSIE_CFG, G1_MINIMAL_CFG, H1_MINIMAL_CFG, UNITREE_A1_CFG,
                      UNITREE_GO1_CFG, UNITREE_GO2_CFG)
from ..envs import mdp
from ..envs.cfg import (ContactSensorCfg, GridPatternCfg, ManagerBasedRLEnvCfg, RayCasterCfg, SceneCfg, SimCfg,
                        TerrainImporterCfg)
from ..managers import CurriculumTermCfg as CurrTerm
from ..managers import EventTermCfg as EventTerm
from ..managers import ObservationGroupCfg as ObsGroup
from ..managers import ObservationTermCfg as ObsTerm
from ..managers import RewardTermCfg as RewTerm
from ..managers import SceneEntityCfg
from ..managers import TerminationTermCfg as DoneTerm
from ..managers import UniformNoiseCfg as Unoise
from ..terrains import ROUGH_TERRAINS_CFG


class _Cfg:
    """Attribute container; terms are listed in insertion order (as the reference iterates cfg.__dict__)."""

    def terms(self):
        return [(k, v) for k, v in self.__dict__.items() if v is not None and not k.startswith("_")]


def _scene(robot):
    return SceneCfg(
        num_envs=4096, env_spacing=2.5, robot=robot,
        terrain=TerrainImporterCfg(terrain_type="generator", terrain_generator=copy.deepcopy(ROUGH_TERRAINS_CFG),
                                   max_init_terrain_level=5),
        height_scanner=RayCasterCfg(prim_path="{ENV_REGEX_NS}/Robot/base", offset=RayCasterCfg.OffsetCfg(pos=(0.0, 0.0, 20.0)),
                                    attach_yaw_only=True, pattern_cfg=GridPatternCfg(resolution=0.1, size=[1.6, 1.0])),
        contact_forces=ContactSensorCfg(prim_path="{ENV_REGEX_NS}/Robot/.*", history_length=3, track_air_time=True))


def _commands():
    c = _Cfg()
    c.base_velocity = mdp.UniformVelocityCommandCfg(
        asset_name="robot", resampling_time_range=(10.0, 10.0), rel_standing_envs=0.02, rel_heading_envs=1.0,
        heading_command=True, heading_control_stiffness=0.5,
        ranges=mdp.UniformVelocityCommandCfg.Ranges(lin_vel_x=(-1.0, 1.0), lin_vel_y=(-1.0, 1.0), ang_vel_z=(-1.0, 1.0),
                                                    heading=(-math.pi, math.pi)))
    return c


def _actions():
    c = _Cfg()
    c.joint_pos = mdp.JointPositionActionCfg(asset_name="robot", joint_names=[".*"], scale=0.5, use_default_offset=True)
    return c


def _observations():
    c = _Cfg()
    p = _Cfg()
    p.base_lin_vel = ObsTerm(func=mdp.base_lin_vel, noise=Unoise(n_min=-0.1, n_max=0.1))
    p.base_ang_vel = ObsTerm(func=mdp.base_ang_vel, noise=Unoise(n_min=-0.2, n_max=0.2))
    p.projected_gravity = ObsTerm(func=mdp.projected_gravity, noise=Unoise(n_min=-0.05, n_max=0.05))
    p.velocity_commands = ObsTerm(func=mdp.generated_commands, params={"command_name": "base_velocity"})
    p.joint_pos = ObsTerm(func=mdp.joint_pos_rel, noise=Unoise(n_min=-0.01, n_max=0.01))
    p.joint_vel = ObsTerm(func=mdp.joint_vel_rel, noise=Unoise(n_min=-1.5, n_max=1.5))
    p.actions = ObsTerm(func=mdp.last_action)
    p.height_scan = ObsTerm(func=mdp.height_scan, params={"sensor_cfg": SceneEntityCfg("height_scanner")},
                            noise=Unoise(n_min=-0.1, n_max=0.1), clip=(-1.0, 1.0))
    p._group = ObsGroup(enable_corruption=True, concatenate_terms=True)
    c.policy = p
    return c


def _events():
    c = _Cfg()
    c.physics_material = EventTerm(func=mdp.randomize_rigid_body_material, mode="startup",
                                   params={"asset_cfg": SceneEntityCfg("robot", body_names=".*"),
                                           "static_friction_range": (0.8, 0.8), "dynamic_friction_range": (0.6, 0.6),
                                           "restitution_range": (0.0, 0.0), "num_buckets": 64})
    c.add_base_mass = EventTerm(func=mdp.randomize_rigid_body_mass, mode="startup",
                                params={"asset_cfg": SceneEntityCfg("robot", body_names="base"),
                                        "mass_distribution_params": (-5.0, 5.0), "operation": "add"})
    c.base_external_force_torque = EventTerm(func=mdp.apply_external_force_torque, mode="reset",
                                             params={"asset_cfg": SceneEntityCfg("robot", body_names="base"),
                                                     "force_range": (0.0, 0.0), "torque_range": (-0.0, 0.0)})
    c.reset_base = EventTerm(func=mdp.reset_root_state_uniform, mode="reset", params={
        "pose_range": {"x": (-0.5, 0.5), "y": (-0.5, 0.5), "yaw": (-3.14, 3.14)},
        "velocity_range": {k: (-0.5, 0.5) for k in ("x", "y", "z", "roll", "pitch", "yaw")}})
    c.reset_robot_joints = EventTerm(func=mdp.reset_joints_by_scale, mode="reset",
                                     params={"position_range": (0.5, 1.5), "velocity_range": (0.0, 0.0)})
    c.push_robot = EventTerm(func=mdp.push_by_setting_velocity, mode="interval", interval_range_s=(10.0, 15.0),
                             params={"velocity_range": {"x": (-0.5, 0.5), "y": (-0.5, 0.5)}})
    return c


def _rewards():
    c = _Cfg()
    c.track_lin_vel_xy_exp = RewTerm(func=mdp.track_lin_vel_xy_exp, weight=1.0,
                                     params={"command_name": "base_velocity", "std": math.sqrt(0.25)})
    c.track_ang_vel_z_exp = RewTerm(func=mdp.track_ang_vel_z_exp, weight=0.5,
                                    params={"command_name": "base_velocity", "std": math.sqrt(0.25)})
    c.lin_vel_z_l2 = RewTerm(func=mdp.lin_vel_z_l2, weight=-2.0)
    c.ang_vel_xy_l2 = RewTerm(func=mdp.ang_vel_xy_l2, weight=-0.05)
    c.dof_torques_l2 = RewTerm(func=mdp.joint_torques_l2, weight=-1.0e-5)
    c.dof_acc_l2 = RewTerm(func=mdp.joint_acc_l2, weight=-2.5e-7)
    c.action_rate_l2 = RewTerm(func=mdp.action_rate_l2, weight=-0.01)
    c.feet_air_time = RewTerm(func=mdp.feet_air_time, weight=0.125,
                              params={"sensor_cfg": SceneEntityCfg("contact_forces", body_names=".*FOOT"),
                                      "command_name": "base_velocity", "threshold": 0.5})
    c.undesired_contacts = RewTerm(func=mdp.undesired_contacts, weight=-1.0,
                                   params={"sensor_cfg": SceneEntityCfg("contact_forces", body_names=".*THIGH"),
                                           "threshold": 1.0})
    c.flat_orientation_l2 = RewTerm(func=mdp.flat_orientation_l2, weight=0.0)
    c.dof_pos_limits = RewTerm(func=mdp.joint_pos_limits, weight=0.0)
    return c


def _terminations():
    c = _Cfg()
    c.time_out = DoneTerm(func=mdp.time_out, time_out=True)
    c.base_contact = DoneTerm(func=mdp.illegal_contact,
                              params={"sensor_cfg": SceneEntityCfg("contact_forces", body_names="base"), "threshold": 1.0})
    return c


def _curriculum():
    c = _Cfg()
    c.terrain_levels = CurrTerm(func=mdp.terrain_levels_vel)
    return c


def LocomotionVelocityRoughEnvCfg(robot) -> ManagerBasedRLEnvCfg:
    """velocity_env_cfg.py:282-310."""
    cfg = ManagerBasedRLEnvCfg(scene=_scene(robot), observations=_observations(), actions=_actions(),
                               commands=_commands(), rewards=_rewards(), terminations=_terminations(),
                               events=_events(), curriculum=_curriculum(), decimation=4, episode_length_s=20.0,
                               sim=SimCfg(dt=0.005, render_interval=4))
    cfg.scene.height_scanner.update_period = cfg.decimation * cfg.sim.dt
    cfg.scene.contact_forces.update_period = cfg.sim.dt
    return cfg


def AnymalCRoughEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/anymal_c/rough_env_cfg.py:17-24 (Isaac-Velocity-Rough-Anymal-C-v0)."""
    return LocomotionVelocityRoughEnvCfg(ANYMAL_C_CFG.replace())


def _flatten(cfg):
    """config/anymal_c/flat_env_cfg.py:12-32: plane terrain, no height scan, no terrain curriculum."""
    cfg.rewards.flat_orientation_l2.weight = -5.0
    cfg.rewards.dof_torques_l2.weight = -2.5e-5
    cfg.rewards.feet_air_time.weight = 0.5
    cfg.scene.terrain.terrain_type = "plane"
    cfg.scene.terrain.terrain_generator = None
    cfg.scene.height_scanner = None
    cfg.observations.policy.height_scan = None
    cfg.curriculum.terrain_levels = None
    return cfg


def AnymalCFlatEnvCfg() -> ManagerBasedRLEnvCfg:
    """Isaac-Velocity-Flat-Anymal-C-v0 (obs 48)."""
    return _flatten(AnymalCRoughEnvCfg())


def UnitreeGo1RoughEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/go1/rough_env_cfg.py:17-60 (Isaac-Velocity-Rough-Unitree-Go1-v0)."""
    return _unitree_rough(UNITREE_GO1_CFG, "trunk")


def UnitreeA1RoughEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/a1/rough_env_cfg.py:18-61 (Isaac-Velocity-Rough-Unitree-A1-v0): Go1's overrides on the A1 (DC-motor actuators)."""
    return _unitree_rough(UNITREE_A1_CFG, "trunk")


def UnitreeGo2RoughEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/go2/rough_env_cfg.py:18-61 (Isaac-Velocity-Rough-Unitree-Go2-v0): base body "base"."""
    return _unitree_rough(UNITREE_GO2_CFG, "base")


def AnymalBRoughEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/anymal_b/rough_env_cfg.py:17-23."""
    return LocomotionVelocityRoughEnvCfg(ANYMAL_B_CFG.replace())


def AnymalDRoughEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/anymal_d/rough_env_cfg.py:17-23."""
    return LocomotionVelocityRoughEnvCfg(ANYMAL_D_CFG.replace())


def _unitree_flat(cfg):
    """config/{a1,go1,go2}/flat_env_cfg.py:12-30: as `_flatten`, with the Unitree reward weights."""
    cfg = _flatten(cfg)
    cfg.rewards.flat_orientation_l2.weight = -2.5
    cfg.rewards.feet_air_time.weight = 0.25
    cfg.rewards.dof_torques_l2.weight = -0.0002   # the rough value: the Unitree flat cfgs do not touch it
    return cfg


def UnitreeGo1FlatEnvCfg(): return _unitree_flat(UnitreeGo1RoughEnvCfg())      # noqa: E704
def UnitreeA1FlatEnvCfg(): return _unitree_flat(UnitreeA1RoughEnvCfg())        # noqa: E704
def UnitreeGo2FlatEnvCfg(): return _unitree_flat(UnitreeGo2RoughEnvCfg())      # noqa: E704
def AnymalBFlatEnvCfg(): return _flatten(AnymalBRoughEnvCfg())                 # noqa: E704
def AnymalDFlatEnvCfg(): return _flatten(AnymalDRoughEnvCfg())                 # noqa: E704


def play(cfg: ManagerBasedRLEnvCfg) -> ManagerBasedRLEnvCfg:
    """The `*_PLAY` variant every config package defines the same way (e.g. config/anymal_c/rough_env_cfg.py:27-47): 50 envs,
    a 5 x 5 terrain without curriculum, spawn on any level, no observation corruption, no pushes."""
    cfg.scene.num_envs = 50
    cfg.scene.env_spacing = 2.5
    if cfg.scene.terrain is not None:
        cfg.scene.terrain.max_init_terrain_level = None
        gen = getattr(cfg.scene.terrain, "terrain_generator", None)
        if gen is not None:
            gen.num_rows, gen.num_cols, gen.curriculum = 5, 5, False
    cfg.observations.policy._group.enable_corruption = False
    cfg.events.base_external_force_torque = None
    cfg.events.push_robot = None
    return cfg


def _unitree_rough(robot, base_body) -> ManagerBasedRLEnvCfg:
    cfg = LocomotionVelocityRoughEnvCfg(robot.replace())
    cfg.scene.height_scanner.prim_path = "{ENV_REGEX_NS}/Robot/" + base_body
    gen = cfg.scene.terrain.terrain_generator
    gen.sub_terrains["boxes"].grid_height_range = (0.025, 0.1)
    gen.sub_terrains["random_rough"].noise_range = (0.01, 0.06)
    gen.sub_terrains["random_rough"].noise_step = 0.01
    cfg.actions.joint_pos.scale = 0.25
    cfg.events.push_robot = None
    cfg.events.add_base_mass.params["mass_distribution_params"] = (-1.0, 3.0)
    cfg.events.add_base_mass.params["asset_cfg"].body_names = base_body
    cfg.events.base_external_force_torque.params["asset_cfg"].body_names = base_body
    cfg.events.reset_robot_joints.params["position_range"] = (1.0, 1.0)
    cfg.events.reset_base.params = {"pose_range": {"x": (-0.5, 0.5), "y": (-0.5, 0.5), "yaw": (-3.14, 3.14)},
                                    "velocity_range": {k: (0.0, 0.0) for k in ("x", "y", "z", "roll", "pitch", "yaw")}}
    cfg.rewards.feet_air_time.params["sensor_cfg"].body_names = ".*_foot"
    cfg.rewards.feet_air_time.weight = 0.01
    cfg.rewards.undesired_contacts = None
    cfg.rewards.dof_torques_l2.weight = -0.0002
    cfg.rewards.track_lin_vel_xy_exp.weight = 1.5
    cfg.rewards.track_ang_vel_z_exp.weight = 0.75
    cfg.rewards.dof_acc_l2.weight = -2.5e-7
    cfg.terminations.base_contact.params["sensor_cfg"].body_names = base_body
    return cfg


def H1RoughEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/h1/rough_env_cfg.py:22-150 (Isaac-Velocity-Rough-H1-v0)."""
    cfg = LocomotionVelocityRoughEnvCfg(H1_MINIMAL_CFG.replace())
    r = cfg.rewards
    # H1Rewards(RewardsCfg): inherited fields keep their position, new fields are appended (dataclass field order)
    r.lin_vel_z_l2 = None
    r.track_lin_vel_xy_exp = RewTerm(func=mdp.track_lin_vel_xy_yaw_frame_exp, weight=1.0,
                                     params={"command_name": "base_velocity", "std": 0.5})
    r.track_ang_vel_z_exp = RewTerm(func=mdp.track_ang_vel_z_world_exp, weight=1.0,
                                    params={"command_name": "base_velocity", "std": 0.5})
    r.feet_air_time = RewTerm(func=mdp.feet_air_time_positive_biped, weight=0.25,
                              params={"command_name": "base_velocity",
                                      "sensor_cfg": SceneEntityCfg("contact_forces", body_names=".*ankle_link"),
                                      "threshold": 0.4})
    r.dof_pos_limits = RewTerm(func=mdp.joint_pos_limits, weight=-1.0,
                               params={"asset_cfg": SceneEntityCfg("robot", joint_names=".*_ankle")})
    r.termination_penalty = RewTerm(func=mdp.is_terminated, weight=-200.0)
    r.feet_slide = RewTerm(func=mdp.feet_slide, weight=-0.25,
                           params={"sensor_cfg": SceneEntityCfg("contact_forces", body_names=".*ankle_link"),
                                   "asset_cfg": SceneEntityCfg("robot", body_names=".*ankle_link")})
    r.joint_deviation_hip = RewTerm(func=mdp.joint_deviation_l1, weight=-0.2,
                                    params={"asset_cfg": SceneEntityCfg("robot", joint_names=[".*_hip_yaw", ".*_hip_roll"])})
    r.joint_deviation_arms = RewTerm(func=mdp.joint_deviation_l1, weight=-0.2,
                                     params={"asset_cfg": SceneEntityCfg("robot", joint_names=[".*_shoulder_.*", ".*_elbow"])})
    r.joint_deviation_torso = RewTerm(func=mdp.joint_deviation_l1, weight=-0.1,
                                      params={"asset_cfg": SceneEntityCfg("robot", joint_names="torso")})
    cfg.scene.height_scanner.prim_path = "{ENV_REGEX_NS}/Robot/torso_link"
    cfg.events.push_robot = None
    cfg.events.add_base_mass = None
    cfg.events.reset_robot_joints.params["position_range"] = (1.0, 1.0)
    cfg.events.base_external_force_torque.params["asset_cfg"].body_names = [".*torso_link"]
    cfg.events.reset_base.params = {"pose_range": {"x": (-0.5, 0.5), "y": (-0.5, 0.5), "yaw": (-3.14, 3.14)},
                                    "velocity_range": {k: (0.0, 0.0) for k in ("x", "y", "z", "roll", "pitch", "yaw")}}
    r.undesired_contacts = None
    r.flat_orientation_l2.weight = -1.0
    r.dof_torques_l2.weight = 0.0
    r.action_rate_l2.weight = -0.005
    r.dof_acc_l2.weight = -1.25e-7
    cfg.commands.base_velocity.ranges.lin_vel_x = (0.0, 1.0)
    cfg.commands.base_velocity.ranges.lin_vel_y = (0.0, 0.0)
    cfg.commands.base_velocity.ranges.ang_vel_z = (-1.0, 1.0)
    cfg.terminations.base_contact.params["sensor_cfg"].body_names = ".*torso_link"
    return cfg


def _flat_biped(cfg):
    """The terrain part of config/{h1,g1,cassie}/flat_env_cfg.py."""
    cfg.scene.terrain.terrain_type = "plane"
    cfg.scene.terrain.terrain_generator = None
    cfg.scene.height_scanner = None
    cfg.observations.policy.height_scan = None
    cfg.curriculum.terrain_levels = None
    return cfg


def H1FlatEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/h1/flat_env_cfg.py:12-25 (Isaac-Velocity-Flat-H1-v0)."""
    cfg = _flat_biped(H1RoughEnvCfg())
    cfg.rewards.feet_air_time.weight = 1.0
    cfg.rewards.feet_air_time.params["threshold"] = 0.6
    return cfg


_ZERO_VEL = {k: (0.0, 0.0) for k in ("x", "y", "z", "roll", "pitch", "yaw")}


def G1RoughEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/g1/rough_env_cfg.py:18-143 (Isaac-Velocity-Rough-G1-v0): 37 joints, H1's reward structure plus the fingers."""
    cfg = LocomotionVelocityRoughEnvCfg(G1_MINIMAL_CFG.replace())
    r = cfg.rewards
    # G1Rewards(RewardsCfg): overridden fields keep their position, new fields are appended in class-body order
    r.track_lin_vel_xy_exp = RewTerm(func=mdp.track_lin_vel_xy_yaw_frame_exp, weight=1.0,
                                     params={"command_name": "base_velocity", "std": 0.5})
    r.track_ang_vel_z_exp = RewTerm(func=mdp.track_ang_vel_z_world_exp, weight=2.0,
                                    params={"command_name": "base_velocity", "std": 0.5})
    r.feet_air_time = RewTerm(func=mdp.feet_air_time_positive_biped, weight=0.25,
                              params={"command_name": "base_velocity",
                                      "sensor_cfg": SceneEntityCfg("contact_forces", body_names=".*_ankle_roll_link"),
                                      "threshold": 0.4})
    r.dof_pos_limits = RewTerm(func=mdp.joint_pos_limits, weight=-1.0,
                               params={"asset_cfg": SceneEntityCfg("robot", joint_names=[".*_ankle_pitch_joint", ".*_ankle_roll_joint"])})
    r.termination_penalty = RewTerm(func=mdp.is_terminated, weight=-200.0)
    r.feet_slide = RewTerm(func=mdp.feet_slide, weight=-0.1,
                           params={"sensor_cfg": SceneEntityCfg("contact_forces", body_names=".*_ankle_roll_link"),
                                   "asset_cfg": SceneEntityCfg("robot", body_names=".*_ankle_roll_link")})
    r.joint_deviation_hip = RewTerm(func=mdp.joint_deviation_l1, weight=-0.1,
                                    params={"asset_cfg": SceneEntityCfg("robot", joint_names=[".*_hip_yaw_joint", ".*_hip_roll_joint"])})
    r.joint_deviation_arms = RewTerm(func=mdp.joint_deviation_l1, weight=-0.1, params={"asset_cfg": SceneEntityCfg(
        "robot", joint_names=[".*_shoulder_pitch_joint", ".*_shoulder_roll_joint", ".*_shoulder_yaw_joint", ".*_elbow_pitch_joint",
                              ".*_elbow_roll_joint"])})
    r.joint_deviation_fingers = RewTerm(func=mdp.joint_deviation_l1, weight=-0.05, params={"asset_cfg": SceneEntityCfg(
        "robot", joint_names=[".*_five_joint", ".*_three_joint", ".*_six_joint", ".*_four_joint", ".*_zero_joint", ".*_one_joint",
                              ".*_two_joint"])})
    r.joint_deviation_torso = RewTerm(func=mdp.joint_deviation_l1, weight=-0.1,
                                      params={"asset_cfg": SceneEntityCfg("robot", joint_names="torso_joint")})
    cfg.scene.height_scanner.prim_path = "{ENV_REGEX_NS}/Robot/torso_link"
    cfg.events.push_robot = None
    cfg.events.add_base_mass = None
    cfg.events.reset_robot_joints.params["position_range"] = (1.0, 1.0)
    cfg.events.base_external_force_torque.params["asset_cfg"].body_names = ["torso_link"]
    cfg.events.reset_base.params = {"pose_range": {"x": (-0.5, 0.5), "y": (-0.5, 0.5), "yaw": (-3.14, 3.14)},
                                    "velocity_range": dict(_ZERO_VEL)}
    r.lin_vel_z_l2.weight = 0.0
    r.undesired_contacts = None
    r.flat_orientation_l2.weight = -1.0
    r.action_rate_l2.weight = -0.005
    r.dof_acc_l2.weight = -1.25e-7
    r.dof_acc_l2.params["asset_cfg"] = SceneEntityCfg("robot", joint_names=[".*_hip_.*", ".*_knee_joint"])
    r.dof_torques_l2.weight = -1.5e-7
    r.dof_torques_l2.params["asset_cfg"] = SceneEntityCfg("robot", joint_names=[".*_hip_.*", ".*_knee_joint", ".*_ankle_.*"])
    cfg.commands.base_velocity.ranges.lin_vel_x = (0.0, 1.0)
    cfg.commands.base_velocity.ranges.lin_vel_y = (-0.0, 0.0)
    cfg.commands.base_velocity.ranges.ang_vel_z = (-1.0, 1.0)
    cfg.terminations.base_contact.params["sensor_cfg"].body_names = "torso_link"
    return cfg


def G1FlatEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/g1/flat_env_cfg.py:14-42 (Isaac-Velocity-Flat-G1-v0)."""
    cfg = _flat_biped(G1RoughEnvCfg())
    r = cfg.rewards
    r.track_ang_vel_z_exp.weight = 1.0
    r.lin_vel_z_l2.weight = -0.2
    r.action_rate_l2.weight = -0.005
    r.dof_acc_l2.weight = -1.0e-7
    r.feet_air_time.weight = 0.75
    r.feet_air_time.params["threshold"] = 0.4
    r.dof_torques_l2.weight = -2.0e-6
    r.dof_torques_l2.params["asset_cfg"] = SceneEntityCfg("robot", joint_names=[".*_hip_.*", ".*_knee_joint"])
    cfg.commands.base_velocity.ranges.lin_vel_x = (0.0, 1.0)
    cfg.commands.base_velocity.ranges.lin_vel_y = (-0.5, 0.5)
    cfg.commands.base_velocity.ranges.ang_vel_z = (-1.0, 1.0)
    return cfg


def CassieRoughEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/cassie/rough_env_cfg.py:18-83 (Isaac-Velocity-Rough-Cassie-v0)."""
    cfg = LocomotionVelocityRoughEnvCfg(CASSIE_CFG.replace())
    r = cfg.rewards
    r.feet_air_time = RewTerm(func=mdp.feet_air_time_positive_biped, weight=2.5,
                              params={"sensor_cfg": SceneEntityCfg("contact_forces", body_names=".*toe"),
                                      "command_name": "base_velocity", "threshold": 0.3})
    r.dof_pos_limits = RewTerm(func=mdp.joint_pos_limits, weight=-1.0,
                               params={"asset_cfg": SceneEntityCfg("robot", joint_names="toe_joint_.*")})
    r.termination_penalty = RewTerm(func=mdp.is_terminated, weight=-200.0)
    r.joint_deviation_hip = RewTerm(func=mdp.joint_deviation_l1, weight=-0.2,
                                    params={"asset_cfg": SceneEntityCfg("robot", joint_names=["hip_abduction_.*", "hip_rotation_.*"])})
    r.joint_deviation_toes = RewTerm(func=mdp.joint_deviation_l1, weight=-0.2,
                                     params={"asset_cfg": SceneEntityCfg("robot", joint_names=["toe_joint_.*"])})
    cfg.scene.height_scanner.prim_path = "{ENV_REGEX_NS}/Robot/pelvis"
    cfg.actions.joint_pos.scale = 0.5
    cfg.events.push_robot = None
    cfg.events.add_base_mass = None
    cfg.events.reset_robot_joints.params["position_range"] = (1.0, 1.0)
    cfg.events.base_external_force_torque.params["asset_cfg"].body_names = [".*pelvis"]
    cfg.events.reset_base.params = {"pose_range": {"x": (-0.5, 0.5), "y": (-0.5, 0.5), "yaw": (-3.14, 3.14)},
                                    "velocity_range": dict(_ZERO_VEL)}
    cfg.terminations.base_contact.params["sensor_cfg"].body_names = [".*pelvis"]
    r.undesired_contacts = None
    r.dof_torques_l2.weight = -5.0e-6
    r.track_lin_vel_xy_exp.weight = 2.0
    r.track_ang_vel_z_exp.weight = 1.0
    r.action_rate_l2.weight *= 1.5
    r.dof_acc_l2.weight *= 1.5
    return cfg


def CassieFlatEnvCfg() -> ManagerBasedRLEnvCfg:
    """config/cassie/flat_env_cfg.py:12-27 (Isaac-Velocity-Flat-Cassie-v0)."""
    cfg = CassieRoughEnvCfg()
    cfg.rewards.flat_orientation_l2.weight = -2.5
    cfg.rewards.feet_air_time.weight = 5.0
    cfg.rewards.joint_deviation_hip.params["asset_cfg"].joint_names = ["hip_rotation_.*"]
    return _flat_biped(cfg)
