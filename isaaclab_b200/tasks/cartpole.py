"""Isaac-Cartpole-v0 as data: TASKS/manager_based/classic/cartpole/cartpole_env_cfg.py:59-181."""
from __future__ import annotations

import math

from ..assets import CARTPOLE_CFG
from ..envs import mdp
from ..envs.cfg import ManagerBasedRLEnvCfg, SceneCfg, SimCfg
from ..managers import EventTermCfg as EventTerm
from ..managers import ObservationGroupCfg as ObsGroup
from ..managers import ObservationTermCfg as ObsTerm
from ..managers import RewardTermCfg as RewTerm
from ..managers import SceneEntityCfg
from ..managers import TerminationTermCfg as DoneTerm
from .velocity import _Cfg


def CartpoleEnvCfg() -> ManagerBasedRLEnvCfg:
    actions = _Cfg()
    actions.joint_effort = mdp.JointEffortActionCfg(asset_name="robot", joint_names=["slider_to_cart"], scale=100.0)
    obs, pol = _Cfg(), _Cfg()
    pol.joint_pos_rel = ObsTerm(func=mdp.joint_pos_rel)
    pol.joint_vel_rel = ObsTerm(func=mdp.joint_vel_rel)
    pol._group = ObsGroup(enable_corruption=False, concatenate_terms=True)
    obs.policy = pol
    ev = _Cfg()
    ev.reset_cart_position = EventTerm(func=mdp.reset_joints_by_offset, mode="reset", params={
        "asset_cfg": SceneEntityCfg("robot", joint_names=["slider_to_cart"]), "position_range": (-1.0, 1.0),
        "velocity_range": (-0.5, 0.5)})
    ev.reset_pole_position = EventTerm(func=mdp.reset_joints_by_offset, mode="reset", params={
        "asset_cfg": SceneEntityCfg("robot", joint_names=["cart_to_pole"]),
        "position_range": (-0.25 * math.pi, 0.25 * math.pi), "velocity_range": (-0.25 * math.pi, 0.25 * math.pi)})
    rew = _Cfg()
    rew.alive = RewTerm(func=mdp.is_alive, weight=1.0)
    rew.terminating = RewTerm(func=mdp.is_terminated, weight=-2.0)
    rew.pole_pos = RewTerm(func=mdp.joint_pos_target_l2, weight=-1.0,
                           params={"asset_cfg": SceneEntityCfg("robot", joint_names=["cart_to_pole"]), "target": 0.0})
    rew.cart_vel = RewTerm(func=mdp.joint_vel_l1, weight=-0.01,
                           params={"asset_cfg": SceneEntityCfg("robot", joint_names=["slider_to_cart"])})
    rew.pole_vel = RewTerm(func=mdp.joint_vel_l1, weight=-0.005,
                           params={"asset_cfg": SceneEntityCfg("robot", joint_names=["cart_to_pole"])})
    term = _Cfg()
    term.time_out = DoneTerm(func=mdp.time_out, time_out=True)
    term.cart_out_of_bounds = DoneTerm(func=mdp.joint_pos_out_of_manual_limit, params={
        "asset_cfg": SceneEntityCfg("robot", joint_names=["slider_to_cart"]), "bounds": (-3.0, 3.0)})
    return ManagerBasedRLEnvCfg(scene=SceneCfg(num_envs=4096, env_spacing=4.0, robot=CARTPOLE_CFG.replace()),
                                observations=obs, actions=actions, commands=None, rewards=rew, terminations=term,
                                events=ev, curriculum=None, decimation=2, episode_length_s=5,
                                sim=SimCfg(dt=1 / 120, render_interval=2))
