"""Actuator cfg classes with the reference's names and fields (LAB/actuators/actuator_cfg.py): ImplicitActuatorCfg,
IdealPDActuatorCfg, DCMotorCfg, DelayedPDActuatorCfg, RemotizedPDActuatorCfg, ActuatorNetLSTMCfg, ActuatorNetMLPCfg.  The models themselves are fused into
`substep_pre` (csrc/mdp.cu); `network_file` may be a TorchScript path or an in-memory module."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any


@dataclass
class ActuatorBaseCfg:
    joint_names_expr: list[str] = None
    effort_limit: dict | float | None = None
    velocity_limit: dict | float | None = None
    stiffness: dict | float | None = None
    damping: dict | float | None = None
    armature: dict | float | None = None
    friction: dict | float | None = None
    kind: str = ""


@dataclass
class ImplicitActuatorCfg(ActuatorBaseCfg):
    kind: str = "implicit"


@dataclass
class IdealPDActuatorCfg(ActuatorBaseCfg):
    kind: str = "ideal_pd"


@dataclass
class DCMotorCfg(IdealPDActuatorCfg):
    saturation_effort: float = None
    kind: str = "dc_motor"


@dataclass
class DelayedPDActuatorCfg(IdealPDActuatorCfg):
    """IdealPD behind a per-env command delay of min_delay..max_delay physics steps, re-drawn at reset
    (LAB/actuators/actuator_cfg.py:216-228, actuator_pd.py:289-347)."""
    min_delay: int = 0
    max_delay: int = 0
    kind: str = "delayed_pd"


@dataclass
class RemotizedPDActuatorCfg(DelayedPDActuatorCfg):
    """DelayedPD whose output torque is clamped to an angle-dependent limit interpolated from `joint_parameter_lookup`
    [n, 3] = (joint angle, transmission ratio, max torque); the box effort / velocity limits are removed
    (LAB/actuators/actuator_cfg.py:261-277, actuator_pd.py:349-416)."""
    joint_parameter_lookup: list = None
    kind: str = "remotized_pd"


@dataclass
class ActuatorNetLSTMCfg(DCMotorCfg):
    network_file: Any = None
    kind: str = "lstm"


@dataclass
class ActuatorNetMLPCfg(DCMotorCfg):
    network_file: Any = None
    pos_scale: float = None
    vel_scale: float = None
    torque_scale: float = None
    input_order: str = "pos_vel"
    input_idx: list[int] = None
    kind: str = "mlp"
