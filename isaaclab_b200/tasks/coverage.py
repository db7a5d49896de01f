"""`B200-Opcode-Coverage-v0`: a synthetic manager cfg (not a reference task) that exercises the fused term opcodes, actuator
kinds and cfg options none of the five BASELINE tasks use -- so that each of them is checked against the UNMODIFIED
reference too (tests/golden/gen_mdp.py builds the twin of this cfg from the reference's own functions and classes:
LAB/envs/mdp/rewards.py:245-283, terminations.py:60-88, observations.py:33,104,140, LAB/actuators/actuator_pd.py:184-286).

Base: the flat Anymal-C velocity task.  Changes: IdealPD + stand-alone DCMotor actuator groups instead of the LSTM net;
action clip; rewards contact_forces / action_l2 / joint_vel_l2 / is_alive / is_terminated; terminations
root_height_below_minimum / bad_orientation; observations base_pos_z / joint_pos / joint_vel with scale and clip;
reset events gated by min_step_count_between_reset and an interval push on one global timer (is_global_time); a DelayedPD
actuator group; observation history (with and without noise) and the scale / clip / bias modifiers."""
from __future__ import annotations

from ..actuators import DCMotorCfg, DelayedPDActuatorCfg, IdealPDActuatorCfg, RemotizedPDActuatorCfg
from ..envs import mdp
from ..managers import ObservationTermCfg as ObsTerm
from ..managers import RewardTermCfg as RewTerm
from ..managers import SceneEntityCfg
from ..managers import TerminationTermCfg as DoneTerm
from ..managers import UniformNoiseCfg as Unoise
from .velocity import AnymalCFlatEnvCfg

# shared with the reference-side twin in tests/golden/gen_mdp.py
ACTUATORS = {
    "hips": dict(kind="ideal_pd", joint_names_expr=[".*F_HAA"], stiffness=40.0, damping=2.0, effort_limit=30.0, velocity_limit=7.5),
    "hips_hind": dict(kind="delayed_pd", joint_names_expr=[".*H_HAA"], stiffness=40.0, damping=2.0, effort_limit=30.0,
                      velocity_limit=7.5, min_delay=1, max_delay=3),
    "legs": dict(kind="dc_motor", joint_names_expr=[".*HFE", ".*F_KFE"], stiffness={".*HFE": 60.0, ".*KFE": 50.0}, damping=3.0,
                 effort_limit=80.0, velocity_limit=7.5, saturation_effort=120.0),
    # angle-dependent torque limit: (joint angle, transmission ratio, max torque) samples; the targets saturate it often
    "knees_hind": dict(kind="remotized_pd", joint_names_expr=[".*H_KFE"], stiffness=50.0, damping=3.0, min_delay=0, max_delay=2,
                       joint_parameter_lookup=[[-1.5, 1.0, 6.0], [-0.5, 1.1, 14.0], [0.4, 1.2, 25.0], [0.9, 1.2, 18.0],
                                               [1.6, 1.0, 9.0]]),
}
ACTION_CLIP = {".*HAA": (-0.8, 0.8), ".*HFE": (-1.5, 1.5), ".*KFE": (-2.0, 1.0)}
REWARDS = {
    "contact_forces": dict(weight=-0.01, params={"sensor_cfg": ("contact_forces", ".*FOOT"), "threshold": 5.0}),
    "action_l2": dict(weight=-0.01, params={}),
    "joint_vel_l2": dict(weight=-1.0e-4, params={"asset_cfg": ("robot", [".*HFE"])}),
    "is_alive": dict(weight=0.5, params={}),
    "is_terminated": dict(weight=-2.0, params={}),
}
TERMINATIONS = {
    "root_height_below_minimum": dict(params={"minimum_height": 0.45}),
    "bad_orientation": dict(params={"limit_angle": 0.2}),
}
OBSERVATIONS = {   # name: (func name, noise range | None, clip | None, scale | None, history_length, modifiers)
    "base_height": ("base_pos_z", None, None, 2.0, 3, None),
    "joint_pos_abs": ("joint_pos", (-0.01, 0.01), (-1.2, 1.2), None, 0,
                      [("scale", {"multiplier": 0.5}), ("clip", {"bounds": (-0.4, None)}), ("bias", {"value": 0.1})]),
    "joint_vel_abs": ("joint_vel", (-0.5, 0.5), (-4.0, 4.0), 0.25, 2, [("clip", {"bounds": (-3.0, 3.0)})]),
}


# EventManager gating (event_manager.py:204-270): reset terms that may only re-trigger after a number of env steps, and an
# interval term on ONE global timer
EVENT_GATES = {"reset_base": 3, "reset_robot_joints": 2}
GLOBAL_PUSH = dict(interval_range_s=(0.03, 0.07), velocity_range={"x": (-0.5, 0.5), "y": (-0.5, 0.5)})


def OpcodeCoverageEnvCfg():
    from ..managers import EventTermCfg
    cfg = AnymalCFlatEnvCfg()
    for name, m in EVENT_GATES.items():
        getattr(cfg.events, name).min_step_count_between_reset = m
    cfg.events.push_all = EventTermCfg(func=mdp.push_by_setting_velocity, mode="interval", is_global_time=True,
                                       interval_range_s=GLOBAL_PUSH["interval_range_s"],
                                       params={"velocity_range": dict(GLOBAL_PUSH["velocity_range"])})
    acts = {}
    for name, a in ACTUATORS.items():
        a = dict(a)
        cls = {"ideal_pd": IdealPDActuatorCfg, "dc_motor": DCMotorCfg, "delayed_pd": DelayedPDActuatorCfg,
               "remotized_pd": RemotizedPDActuatorCfg}[a.pop("kind")]
        acts[name] = cls(**a)
    cfg.scene.robot.actuators = acts
    cfg.actions.joint_pos.clip = dict(ACTION_CLIP)
    for name, r in REWARDS.items():
        params = {}
        for k, v in r["params"].items():
            if k == "sensor_cfg":
                v = SceneEntityCfg(v[0], body_names=v[1])
            elif k == "asset_cfg":
                v = SceneEntityCfg(v[0], joint_names=v[1])
            params[k] = v
        setattr(cfg.rewards, "cov_" + name, RewTerm(func=getattr(mdp, name), weight=r["weight"], params=params))
    for name, t in TERMINATIONS.items():
        setattr(cfg.terminations, "cov_" + name, DoneTerm(func=getattr(mdp, name), params=dict(t["params"])))
    from ..utils import modifiers
    for name, (fn, nz, clip, scale, hist, mods) in OBSERVATIONS.items():
        mods = [modifiers.ModifierCfg(func=getattr(modifiers, m), params=dict(pr)) for m, pr in mods] if mods else None
        setattr(cfg.observations.policy, name,
                ObsTerm(func=getattr(mdp, fn), noise=Unoise(n_min=nz[0], n_max=nz[1]) if nz else None, clip=clip, scale=scale,
                        history_length=hist, modifiers=mods))
    return cfg
