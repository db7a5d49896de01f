"""Robot descriptions as data: the `ArticulationCfg`s of ASSETS/robots/{anymal,unitree,cartpole}.py (init state, actuator
groups, soft-limit factor) plus what the reference reads from the USD file at run time and therefore is NOT in its
tree -- joint names/order, body names/order, joint position limits.  Those are stated here explicitly (PhysX
breadth-first ordering of the public robot models); when a recorded Isaac Sim state is replayed, the recording's
own name lists replace them.
"""
from __future__ import annotations

import copy
from dataclasses import dataclass, field

import torch

from .actuators import ActuatorNetLSTMCfg, ActuatorNetMLPCfg, DCMotorCfg, ImplicitActuatorCfg


@dataclass
class ArticulationCfg:
    joint_names: list[str] = None
    body_names: list[str] = None
    init_pos: tuple = (0.0, 0.0, 0.0)
    init_rot: tuple = (1.0, 0.0, 0.0, 0.0)
    init_lin_vel: tuple = (0.0, 0.0, 0.0)
    init_ang_vel: tuple = (0.0, 0.0, 0.0)
    joint_pos: dict = field(default_factory=lambda: {".*": 0.0})
    joint_vel: dict = field(default_factory=lambda: {".*": 0.0})
    joint_pos_limits: dict = field(default_factory=lambda: {".*": (-3.14, 3.14)})
    soft_joint_pos_limit_factor: float = 1.0
    actuators: dict = field(default_factory=dict)

    def replace(self, **kw):
        c = copy.deepcopy(self)
        for k, v in kw.items():
            if hasattr(c, k):
                setattr(c, k, v)
        return c

    copy = replace


def synthetic_lstm_net(hidden: int = 8, num_layers: int = 2, seed: int = 0) -> dict:
    """Weights with the documented structure of the ANYdrive actuator net (`network.lstm` = nn.LSTM(2, H, L) + a
    Linear(H, 1) head; LAB/actuators/actuator_net.py:53-54).  The real file lives on the Nucleus server
    (ASSETS/robots/anymal.py:47) and is not in the reference tree."""
    g = torch.Generator().manual_seed(seed)
    k = 1.0 / hidden ** 0.5
    u = lambda *s: (torch.rand(*s, generator=g) * 2 - 1) * k  # noqa: E731  (nn.LSTM default init)
    net = {"num_layers": num_layers, "hidden": hidden}
    for l in range(num_layers):
        net[f"weight_ih_l{l}"] = u(4 * hidden, 2 if l == 0 else hidden)
        net[f"weight_hh_l{l}"] = u(4 * hidden, hidden)
        net[f"bias_ih_l{l}"] = u(4 * hidden)
        net[f"bias_hh_l{l}"] = u(4 * hidden)
    net["head_weight"] = u(1, hidden) * 20.0
    net["head_bias"] = u(1)
    return net


def synthetic_mlp_net(cfg: ActuatorNetMLPCfg, hidden=(32, 32, 32), activation="softsign", seed: int = 0) -> dict:
    """MLP 2*len(input_idx) -> 32 -> 32 -> 32 -> 1 with softsign, the walk-these-ways Go1 actuator model the reference
    cites (ASSETS/robots/unitree.py:30-46); weights synthetic for the same reason as above."""
    g = torch.Generator().manual_seed(seed)
    dims = [2 * len(cfg.input_idx), *hidden, 1]
    layers = []
    for i in range(len(dims) - 1):
        k = 1.0 / dims[i] ** 0.5
        layers.append(((torch.rand(dims[i + 1], dims[i], generator=g) * 2 - 1) * k,
                       (torch.rand(dims[i + 1], generator=g) * 2 - 1) * k))
    layers[-1] = (layers[-1][0] * 10.0, layers[-1][1])
    return {"layers": layers, "activation": activation, "input_idx": list(cfg.input_idx), "input_order": cfg.input_order,
            "pos_scale": cfg.pos_scale, "vel_scale": cfg.vel_scale, "torque_scale": cfg.torque_scale}


def load_lstm_net(path) -> dict:
    """Weights of a TorchScript ANYdrive-style net (`.lstm` attribute + one Linear head)."""
    m = torch.jit.load(path, map_location="cpu").eval()
    sd = m.lstm.state_dict()
    net = {"num_layers": len(sd) // 4, "hidden": sd["weight_hh_l0"].shape[1]}
    net.update({k: v.float() for k, v in sd.items()})
    rest = {k: v for k, v in m.state_dict().items() if not k.startswith("lstm.")}
    w = [v for k, v in rest.items() if k.endswith("weight")]
    b = [v for k, v in rest.items() if k.endswith("bias")]
    if len(w) != 1 or len(b) != 1:
        raise ValueError("expected exactly one Linear head after the LSTM")
    net["head_weight"], net["head_bias"] = w[0].float(), b[0].float()
    return net


_ANYMAL_LEGS = ["LF", "LH", "RF", "RH"]
ANYDRIVE_3_LSTM_ACTUATOR_CFG = ActuatorNetLSTMCfg(joint_names_expr=[".*HAA", ".*HFE", ".*KFE"], network_file=None,
                                                  saturation_effort=120.0, effort_limit=80.0, velocity_limit=7.5)
ANYDRIVE_3_SIMPLE_ACTUATOR_CFG = DCMotorCfg(joint_names_expr=[".*HAA", ".*HFE", ".*KFE"], saturation_effort=120.0,
                                            effort_limit=80.0, velocity_limit=7.5, stiffness={".*": 40.0},
                                            damping={".*": 5.0})
ANYMAL_C_CFG = ArticulationCfg(  # ASSETS/robots/anymal.py:93-124
    joint_names=[f"{l}_{j}" for j in ("HAA", "HFE", "KFE") for l in _ANYMAL_LEGS],
    body_names=["base"] + [f"{l}_{b}" for b in ("HIP", "THIGH", "SHANK", "FOOT") for l in _ANYMAL_LEGS],
    init_pos=(0.0, 0.0, 0.6),
    joint_pos={".*HAA": 0.0, ".*F_HFE": 0.4, ".*H_HFE": -0.4, ".*F_KFE": -0.8, ".*H_KFE": 0.8},
    joint_pos_limits={".*HAA": (-0.72, 0.49), ".*HFE": (-9.42, 9.42), ".*KFE": (-9.42, 9.42)},
    actuators={"legs": ANYDRIVE_3_LSTM_ACTUATOR_CFG}, soft_joint_pos_limit_factor=0.95)

_GO1_LEGS = ["FL", "FR", "RL", "RR"]
GO1_ACTUATOR_CFG = ActuatorNetMLPCfg(  # ASSETS/robots/unitree.py:30-41
    joint_names_expr=[".*_hip_joint", ".*_thigh_joint", ".*_calf_joint"], network_file=None, pos_scale=-1.0,
    vel_scale=1.0, torque_scale=1.0, input_order="pos_vel", input_idx=[0, 1, 2], effort_limit=23.7,
    velocity_limit=30.0, saturation_effort=23.7)
UNITREE_GO1_CFG = ArticulationCfg(  # unitree.py:102-134
    joint_names=[f"{l}_{j}_joint" for j in ("hip", "thigh", "calf") for l in _GO1_LEGS],
    body_names=["trunk"] + [f"{l}_{b}" for b in ("hip", "thigh", "calf", "foot") for l in _GO1_LEGS],
    init_pos=(0.0, 0.0, 0.4),
    joint_pos={".*L_hip_joint": 0.1, ".*R_hip_joint": -0.1, "F[L,R]_thigh_joint": 0.8, "R[L,R]_thigh_joint": 1.0,
               ".*_calf_joint": -1.5},
    joint_pos_limits={".*_hip_joint": (-0.863, 0.863), ".*_thigh_joint": (-0.686, 4.501), ".*_calf_joint": (-2.818, -0.888)},
    actuators={"base_legs": GO1_ACTUATOR_CFG}, soft_joint_pos_limit_factor=0.9)

# ANYmal-B / ANYmal-D: the same leg layout, defaults and ANYdrive 3.0 actuator net as ANYmal-C (ASSETS/robots/anymal.py:59-160);
# the three differ in their USD geometry and masses, which live on the simulator side of the boundary
ANYMAL_B_CFG = ANYMAL_C_CFG.copy()
ANYMAL_D_CFG = ANYMAL_C_CFG.copy()


def _unitree_quadruped(base_body, init_z, effort, velocity):
    """A1 / Go2 (ASSETS/robots/unitree.py:49-97,138-176): Go1's joint layout driven by a DC-motor model."""
    return ArticulationCfg(
        joint_names=[f"{l}_{j}_joint" for j in ("hip", "thigh", "calf") for l in _GO1_LEGS],
        body_names=[base_body] + [f"{l}_{b}" for b in ("hip", "thigh", "calf", "foot") for l in _GO1_LEGS],
        init_pos=(0.0, 0.0, init_z),
        joint_pos={".*L_hip_joint": 0.1, ".*R_hip_joint": -0.1, "F[L,R]_thigh_joint": 0.8, "R[L,R]_thigh_joint": 1.0,
                   ".*_calf_joint": -1.5},
        joint_pos_limits={".*_hip_joint": (-0.863, 0.863), ".*_thigh_joint": (-0.686, 4.501), ".*_calf_joint": (-2.818, -0.888)},
        actuators={"base_legs": DCMotorCfg(joint_names_expr=[".*_hip_joint", ".*_thigh_joint", ".*_calf_joint"],
                                           effort_limit=effort, saturation_effort=effort, velocity_limit=velocity,
                                           stiffness=25.0, damping=0.5, friction=0.0)},
        soft_joint_pos_limit_factor=0.9)


UNITREE_A1_CFG = _unitree_quadruped("trunk", 0.42, 33.5, 21.0)
UNITREE_GO2_CFG = _unitree_quadruped("base", 0.4, 23.5, 30.0)

_H1_JOINTS = ["left_hip_yaw", "right_hip_yaw", "torso", "left_hip_roll", "right_hip_roll", "left_shoulder_pitch",
              "right_shoulder_pitch", "left_hip_pitch", "right_hip_pitch", "left_shoulder_roll", "right_shoulder_roll",
              "left_knee", "right_knee", "left_shoulder_yaw", "right_shoulder_yaw", "left_ankle", "right_ankle",
              "left_elbow", "right_elbow"]
H1_CFG = ArticulationCfg(  # unitree.py:182-261
    joint_names=_H1_JOINTS, body_names=["pelvis"] + [f"{j}_link" for j in _H1_JOINTS], init_pos=(0.0, 0.0, 1.05),
    joint_pos={".*_hip_yaw": 0.0, ".*_hip_roll": 0.0, ".*_hip_pitch": -0.28, ".*_knee": 0.79, ".*_ankle": -0.52,
               "torso": 0.0, ".*_shoulder_pitch": 0.28, ".*_shoulder_roll": 0.0, ".*_shoulder_yaw": 0.0, ".*_elbow": 0.52},
    joint_pos_limits={".*_hip_yaw": (-0.43, 0.43), ".*_hip_roll": (-0.43, 0.43), ".*_hip_pitch": (-3.14, 2.53),
                      ".*_knee": (-0.26, 2.05), ".*_ankle": (-0.87, 0.52), "torso": (-2.35, 2.35),
                      ".*_shoulder_pitch": (-2.87, 2.87), ".*_shoulder_roll": (-3.11, 3.11),
                      ".*_shoulder_yaw": (-4.45, 4.45), ".*_elbow": (-1.25, 2.61)},
    soft_joint_pos_limit_factor=0.9,
    actuators={
        "legs": ImplicitActuatorCfg(joint_names_expr=[".*_hip_yaw", ".*_hip_roll", ".*_hip_pitch", ".*_knee", "torso"],
                                    effort_limit=300, velocity_limit=100.0,
                                    stiffness={".*_hip_yaw": 150.0, ".*_hip_roll": 150.0, ".*_hip_pitch": 200.0,
                                               ".*_knee": 200.0, "torso": 200.0},
                                    damping={".*_hip_yaw": 5.0, ".*_hip_roll": 5.0, ".*_hip_pitch": 5.0, ".*_knee": 5.0,
                                             "torso": 5.0}),
        "feet": ImplicitActuatorCfg(joint_names_expr=[".*_ankle"], effort_limit=100, velocity_limit=100.0,
                                    stiffness={".*_ankle": 20.0}, damping={".*_ankle": 4.0}),
        "arms": ImplicitActuatorCfg(joint_names_expr=[".*_shoulder_pitch", ".*_shoulder_roll", ".*_shoulder_yaw", ".*_elbow"],
                                    effort_limit=300, velocity_limit=100.0,
                                    stiffness={".*_shoulder_pitch": 40.0, ".*_shoulder_roll": 40.0,
                                               ".*_shoulder_yaw": 40.0, ".*_elbow": 40.0},
                                    damping={".*_shoulder_pitch": 10.0, ".*_shoulder_roll": 10.0,
                                             ".*_shoulder_yaw": 10.0, ".*_elbow": 10.0}),
    })
H1_MINIMAL_CFG = H1_CFG.copy()

# Unitree G1 (ASSETS/robots/unitree.py:273-381): 37 joints -- 2 x 6 leg, torso, 2 x 5 arm, 2 x 7 hand.  The joint / body
# ORDER of the real asset is PhysX's traversal of its USD; here it is a fixed listing (names as the cfg's regexes need them)
_G1_SIDES = ("left", "right")
_G1_JOINTS = ([f"{s}_{j}_joint" for j in ("hip_pitch", "hip_roll", "hip_yaw", "knee", "ankle_pitch", "ankle_roll") for s in _G1_SIDES]
              + ["torso_joint"]
              + [f"{s}_{j}_joint" for j in ("shoulder_pitch", "shoulder_roll", "shoulder_yaw", "elbow_pitch", "elbow_roll") for s in _G1_SIDES]
              + [f"{s}_{j}_joint" for j in ("zero", "one", "two", "three", "four", "five", "six") for s in _G1_SIDES])
_G1_ARM_JOINTS = [".*_shoulder_pitch_joint", ".*_shoulder_roll_joint", ".*_shoulder_yaw_joint", ".*_elbow_pitch_joint",
                  ".*_elbow_roll_joint", ".*_five_joint", ".*_three_joint", ".*_six_joint", ".*_four_joint", ".*_zero_joint",
                  ".*_one_joint", ".*_two_joint"]
G1_CFG = ArticulationCfg(
    joint_names=_G1_JOINTS, body_names=["pelvis"] + [j[:-len("_joint")] + "_link" for j in _G1_JOINTS], init_pos=(0.0, 0.0, 0.74),
    joint_pos={".*_hip_pitch_joint": -0.20, ".*_knee_joint": 0.42, ".*_ankle_pitch_joint": -0.23, ".*_elbow_pitch_joint": 0.87,
               "left_shoulder_roll_joint": 0.16, "left_shoulder_pitch_joint": 0.35, "right_shoulder_roll_joint": -0.16,
               "right_shoulder_pitch_joint": 0.35, "left_one_joint": 1.0, "right_one_joint": -1.0, "left_two_joint": 0.52,
               "right_two_joint": -0.52},
    joint_pos_limits={".*_hip_pitch_joint": (-2.35, 3.05), ".*_hip_roll_joint": (-0.26, 2.53), ".*_hip_yaw_joint": (-2.75, 2.75),
                      ".*_knee_joint": (-0.33, 2.54), ".*_ankle_pitch_joint": (-0.68, 0.73), ".*_ankle_roll_joint": (-0.26, 0.26),
                      "torso_joint": (-2.62, 2.62), ".*_shoulder_pitch_joint": (-2.97, 2.79), ".*_shoulder_roll_joint": (-2.25, 2.25),
                      ".*_shoulder_yaw_joint": (-2.62, 2.62), ".*_elbow_pitch_joint": (-0.23, 3.42),
                      ".*_elbow_roll_joint": (-2.09, 2.09), ".*_(zero|one|two|three|four|five|six)_joint": (-1.84, 1.84)},
    soft_joint_pos_limit_factor=0.9,
    actuators={
        "legs": ImplicitActuatorCfg(joint_names_expr=[".*_hip_yaw_joint", ".*_hip_roll_joint", ".*_hip_pitch_joint", ".*_knee_joint",
                                                      "torso_joint"], effort_limit=300, velocity_limit=100.0,
                                    stiffness={".*_hip_yaw_joint": 150.0, ".*_hip_roll_joint": 150.0, ".*_hip_pitch_joint": 200.0,
                                               ".*_knee_joint": 200.0, "torso_joint": 200.0},
                                    damping={".*_hip_yaw_joint": 5.0, ".*_hip_roll_joint": 5.0, ".*_hip_pitch_joint": 5.0,
                                             ".*_knee_joint": 5.0, "torso_joint": 5.0},
                                    armature={".*_hip_.*": 0.01, ".*_knee_joint": 0.01, "torso_joint": 0.01}),
        "feet": ImplicitActuatorCfg(effort_limit=20, joint_names_expr=[".*_ankle_pitch_joint", ".*_ankle_roll_joint"],
                                    stiffness=20.0, damping=2.0, armature=0.01),
        "arms": ImplicitActuatorCfg(joint_names_expr=list(_G1_ARM_JOINTS), effort_limit=300, velocity_limit=100.0, stiffness=40.0,
                                    damping=10.0,
                                    armature={".*_shoulder_.*": 0.01, ".*_elbow_.*": 0.01, ".*_five_joint": 0.001,
                                              ".*_three_joint": 0.001, ".*_six_joint": 0.001, ".*_four_joint": 0.001,
                                              ".*_zero_joint": 0.001, ".*_one_joint": 0.001, ".*_two_joint": 0.001}),
    })
G1_MINIMAL_CFG = G1_CFG.copy()

# Agility Cassie (ASSETS/robots/cassie.py:24-92): the 12 joints the cfg names
_CASSIE_JOINTS = [f"{j}_{s}" for j in ("hip_abduction", "hip_rotation", "hip_flexion", "thigh_joint", "ankle_joint", "toe_joint")
                  for s in ("left", "right")]
CASSIE_CFG = ArticulationCfg(
    joint_names=_CASSIE_JOINTS,
    body_names=["pelvis"] + [f"{s}_{b}" for b in ("pelvis_abduction", "pelvis_rotation", "hip", "thigh", "shin", "toe")
                             for s in ("left", "right")],
    init_pos=(0.0, 0.0, 0.9),
    joint_pos={"hip_abduction_left": 0.1, "hip_rotation_left": 0.0, "hip_flexion_left": 1.0, "thigh_joint_left": -1.8,
               "ankle_joint_left": 1.57, "toe_joint_left": -1.57, "hip_abduction_right": -0.1, "hip_rotation_right": 0.0,
               "hip_flexion_right": 1.0, "thigh_joint_right": -1.8, "ankle_joint_right": 1.57, "toe_joint_right": -1.57},
    joint_pos_limits={"hip_abduction_.*": (-0.26, 0.39), "hip_rotation_.*": (-0.39, 0.39), "hip_flexion_.*": (-0.87, 1.40),
                      "thigh_joint_.*": (-2.86, -0.65), "ankle_joint_.*": (0.87, 2.97), "toe_joint_.*": (-2.44, -0.52)},
    soft_joint_pos_limit_factor=0.9,
    actuators={
        "legs": ImplicitActuatorCfg(joint_names_expr=["hip_.*", "thigh_.*", "ankle_.*"], effort_limit=200.0, velocity_limit=10.0,
                                    stiffness={"hip_abduction.*": 100.0, "hip_rotation.*": 100.0, "hip_flexion.*": 200.0,
                                               "thigh_joint.*": 200.0, "ankle_joint.*": 200.0},
                                    damping={"hip_abduction.*": 3.0, "hip_rotation.*": 3.0, "hip_flexion.*": 6.0,
                                             "thigh_joint.*": 6.0, "ankle_joint.*": 6.0}),
        "toes": ImplicitActuatorCfg(joint_names_expr=["toe_.*"], effort_limit=20.0, velocity_limit=10.0,
                                    stiffness={"toe_joint.*": 20.0}, damping={"toe_joint.*": 1.0}),
    })

CARTPOLE_CFG = ArticulationCfg(  # ASSETS/robots/cartpole.py
    joint_names=["slider_to_cart", "cart_to_pole"], body_names=["rail", "cart", "pole"], init_pos=(0.0, 0.0, 2.0),
    joint_pos={"slider_to_cart": 0.0, "cart_to_pole": 0.0},
    joint_pos_limits={"slider_to_cart": (-4.0, 4.0), "cart_to_pole": (-1.0e6, 1.0e6)},
    actuators={
        "cart_actuator": ImplicitActuatorCfg(joint_names_expr=["slider_to_cart"], effort_limit=400.0,
                                             velocity_limit=100.0, stiffness=0.0, damping=10.0),
        "pole_actuator": ImplicitActuatorCfg(joint_names_expr=["cart_to_pole"], effort_limit=400.0,
                                             velocity_limit=100.0, stiffness=0.0, damping=0.0),
    })
