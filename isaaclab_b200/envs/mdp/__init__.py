"""The MDP term library, by the reference's names (`isaaclab.envs.mdp.*`, the locomotion-velocity `mdp` package and
the cartpole `mdp` package).  Each function is the IDENTITY of a term that has a fused opcode in libb200lab: task cfgs
reference them exactly as they reference the reference's functions (`func=mdp.track_lin_vel_xy_exp`), and
`isaaclab_b200.envs.compile` lowers recognised identities to opcodes.  Calling one directly evaluates that single
term on the env's current state through `ManagerBasedRLEnv.evaluate_term` (same kernels, one-term program).

Citations give the reference definition each identity stands for.
LAB = source/isaaclab/isaaclab, TASKS = source/isaaclab_tasks/isaaclab_tasks.
"""
from __future__ import annotations

from dataclasses import dataclass, field

from ...managers import (ActionTermCfg, CommandTermCfg, SceneEntityCfg)

_REGISTRY = {}


def _term(kind, op, ref, defaults=None):
    def fn(env, *args, **params):
        return env.evaluate_term(fn, *args, **params)

    fn.__name__ = fn.__qualname__ = op
    fn.__doc__ = f"Fused term `{op}` ({kind}); reference: {ref}"
    fn._b200_kind, fn._b200_op, fn._b200_defaults = kind, op, dict(defaults or {})
    _REGISTRY[op] = fn
    return fn


# observations -- LAB/envs/mdp/observations.py
base_pos_z = _term("obs", "base_pos_z", "LAB/envs/mdp/observations.py:33")
base_lin_vel = _term("obs", "base_lin_vel", "observations.py:40")
base_ang_vel = _term("obs", "base_ang_vel", "observations.py:47")
projected_gravity = _term("obs", "projected_gravity", "observations.py:54")
joint_pos = _term("obs", "joint_pos", "observations.py:104")
joint_pos_rel = _term("obs", "joint_pos_rel", "observations.py:114")
joint_vel = _term("obs", "joint_vel", "observations.py:140")
joint_vel_rel = _term("obs", "joint_vel_rel", "observations.py:150")
height_scan = _term("obs", "height_scan", "observations.py:165", {"offset": 0.5})
last_action = _term("obs", "last_action", "observations.py:512")
generated_commands = _term("obs", "generated_commands", "observations.py:529")
# rewards -- LAB/envs/mdp/rewards.py
is_alive = _term("rew", "is_alive", "LAB/envs/mdp/rewards.py:31")
is_terminated = _term("rew", "is_terminated", "rewards.py:36")
lin_vel_z_l2 = _term("rew", "lin_vel_z_l2", "rewards.py:76")
ang_vel_xy_l2 = _term("rew", "ang_vel_xy_l2", "rewards.py:83")
flat_orientation_l2 = _term("rew", "flat_orientation_l2", "rewards.py:90")
joint_torques_l2 = _term("rew", "joint_torques_l2", "rewards.py:136")
joint_vel_l1 = _term("rew", "joint_vel_l1", "rewards.py:146")
joint_vel_l2 = _term("rew", "joint_vel_l2", "rewards.py:153")
joint_acc_l2 = _term("rew", "joint_acc_l2", "rewards.py:163")
joint_deviation_l1 = _term("rew", "joint_deviation_l1", "rewards.py:173")
joint_pos_limits = _term("rew", "joint_pos_limits", "rewards.py:182")
action_rate_l2 = _term("rew", "action_rate_l2", "rewards.py:245")
action_l2 = _term("rew", "action_l2", "rewards.py:250")
undesired_contacts = _term("rew", "undesired_contacts", "rewards.py:260")
contact_forces = _term("rew", "contact_forces", "rewards.py:271")
track_lin_vel_xy_exp = _term("rew", "track_lin_vel_xy_exp", "rewards.py:287")
track_ang_vel_z_exp = _term("rew", "track_ang_vel_z_exp", "rewards.py:301")
# rewards -- TASKS/manager_based/locomotion/velocity/mdp/rewards.py, TASKS/manager_based/classic/cartpole/mdp/rewards.py
feet_air_time = _term("rew", "feet_air_time", "TASKS/.../velocity/mdp/rewards.py:25")
feet_air_time_positive_biped = _term("rew", "feet_air_time_positive_biped", "velocity/mdp/rewards.py:47")
feet_slide = _term("rew", "feet_slide", "velocity/mdp/rewards.py:69")
track_lin_vel_xy_yaw_frame_exp = _term("rew", "track_lin_vel_xy_yaw_frame_exp", "velocity/mdp/rewards.py:86")
track_ang_vel_z_world_exp = _term("rew", "track_ang_vel_z_world_exp", "velocity/mdp/rewards.py:99")
joint_pos_target_l2 = _term("rew", "joint_pos_target_l2", "TASKS/.../cartpole/mdp/rewards.py:19")
# terminations -- LAB/envs/mdp/terminations.py
time_out = _term("term", "time_out", "LAB/envs/mdp/terminations.py:30")
bad_orientation = _term("term", "bad_orientation", "terminations.py:60")
root_height_below_minimum = _term("term", "root_height_below_minimum", "terminations.py:74")
joint_pos_out_of_manual_limit = _term("term", "joint_pos_out_of_manual_limit", "terminations.py:90")
illegal_contact = _term("term", "illegal_contact", "terminations.py:150")
# events -- LAB/envs/mdp/events.py
reset_root_state_uniform = _term("event", "reset_root_state_uniform", "LAB/envs/mdp/events.py:823")
reset_joints_by_scale = _term("event", "reset_joints_by_scale", "events.py:987")
reset_joints_by_offset = _term("event", "reset_joints_by_offset", "events.py:1020")
push_by_setting_velocity = _term("event", "push_by_setting_velocity", "events.py:795")
apply_external_force_torque = _term("event", "apply_external_force_torque", "events.py:764 (writes wrenches to PhysX)")
randomize_rigid_body_material = _term("event", "randomize_rigid_body_material", "events.py (startup, PhysX property)")
randomize_rigid_body_mass = _term("event", "randomize_rigid_body_mass", "events.py (startup, PhysX property)")
# curriculum -- TASKS/manager_based/locomotion/velocity/mdp/curriculums.py
terrain_levels_vel = _term("curriculum", "terrain_levels_vel", "TASKS/.../velocity/mdp/curriculums.py:26")


# action terms -- LAB/envs/mdp/actions/actions_cfg.py + joint_actions.py
class JointPositionAction:  # identities of the class_type field
    pass


class JointEffortAction:
    pass


@dataclass
class JointActionCfg(ActionTermCfg):
    joint_names: list[str] = None
    scale: float | dict = 1.0
    offset: float | dict = 0.0
    preserve_order: bool = False


@dataclass
class JointPositionActionCfg(JointActionCfg):
    class_type: type = JointPositionAction
    use_default_offset: bool = True


@dataclass
class JointEffortActionCfg(JointActionCfg):
    class_type: type = JointEffortAction


# command terms -- LAB/envs/mdp/commands/commands_cfg.py + velocity_command.py
class UniformVelocityCommand:
    pass


@dataclass
class UniformVelocityCommandCfg(CommandTermCfg):
    @dataclass
    class Ranges:
        lin_vel_x: tuple[float, float] = None
        lin_vel_y: tuple[float, float] = None
        ang_vel_z: tuple[float, float] = None
        heading: tuple[float, float] | None = None

    class_type: type = UniformVelocityCommand
    asset_name: str = "robot"
    heading_command: bool = False
    heading_control_stiffness: float = 1.0
    rel_standing_envs: float = 0.0
    rel_heading_envs: float = 1.0
    ranges: "UniformVelocityCommandCfg.Ranges" = None


__all__ = [n for n in dir() if not n.startswith("_")]
