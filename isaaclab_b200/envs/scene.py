"""What a Python term reads from `env.scene` -- `ArticulationData`, `ContactSensorData`, `RayCasterData` shaped VIEWS of
the buffers the fused kernels own (no copies; derived quantities are computed with torch on access, as
ArticulationData's lazy properties do: LAB/assets/articulation/articulation_data.py:366-619).  Only user (EXTERNAL)
terms touch these; the fused term programs never do."""
from __future__ import annotations

import torch

from ..utils import math as math_utils


class ArticulationDataView:
    def __init__(self, env):
        self._env = env

    # -- simulator state of the current step (the view the last kernels were given)
    @property
    def _sim(self):
        return self._env.sim.last_view()

    @property
    def root_state_w(self):
        return self._sim.root_state_w

    @property
    def root_pos_w(self):
        return self._sim.root_state_w[:, 0:3]

    @property
    def root_quat_w(self):
        return self._sim.root_state_w[:, 3:7]

    @property
    def root_lin_vel_w(self):
        return self._sim.root_state_w[:, 7:10]

    @property
    def root_ang_vel_w(self):
        return self._sim.root_state_w[:, 10:13]

    @property
    def root_lin_vel_b(self):
        return math_utils.quat_rotate_inverse(self.root_quat_w, self.root_lin_vel_w)

    @property
    def root_ang_vel_b(self):
        return math_utils.quat_rotate_inverse(self.root_quat_w, self.root_ang_vel_w)

    @property
    def projected_gravity_b(self):
        g = torch.tensor([0.0, 0.0, -1.0], device=self.root_quat_w.device).repeat(self.root_quat_w.shape[0], 1)
        return math_utils.quat_rotate_inverse(self.root_quat_w, g)

    @property
    def heading_w(self):
        fwd = torch.tensor([1.0, 0.0, 0.0], device=self.root_quat_w.device).repeat(self.root_quat_w.shape[0], 1)
        f = math_utils.quat_apply(self.root_quat_w, fwd)
        return torch.atan2(f[:, 1], f[:, 0])

    @property
    def joint_pos(self):
        return self._sim.joint_pos

    @property
    def joint_vel(self):
        return self._sim.joint_vel

    @property
    def body_lin_vel_w(self):
        return self._sim.body_lin_vel_w

    # -- buffers owned by the fused env
    def __getattr__(self, name):
        f = self.__dict__["_env"].fused
        if name in ("joint_acc", "applied_torque", "computed_torque", "joint_pos_target", "joint_vel_target",
                    "joint_effort_target", "soft_joint_pos_limits", "soft_joint_vel_limits"):
            return getattr(f, name)
        if name in ("default_joint_pos", "default_joint_vel"):
            return getattr(f, name).unsqueeze(0).expand(f.N, -1)
        if name == "default_root_state":
            return f.default_root_state.unsqueeze(0).expand(f.N, -1)
        raise AttributeError(f"ArticulationData view has no '{name}'")


class ArticulationView:
    def __init__(self, env):
        self.data = ArticulationDataView(env)
        self._env = env

    @property
    def joint_names(self):
        return self._env.spec["joint_names"]

    @property
    def body_names(self):
        return self._env.spec["body_names"]

    @property
    def num_joints(self):
        return self._env.fused.J

    @property
    def num_bodies(self):
        return self._env.fused.B

    def find_joints(self, keys, preserve_order=False):
        from ..utils.string import resolve_matching_names
        return resolve_matching_names(keys, self.joint_names, preserve_order)

    def find_bodies(self, keys, preserve_order=False):
        from ..utils.string import resolve_matching_names
        return resolve_matching_names(keys, self.body_names, preserve_order)


class ContactSensorDataView:
    def __init__(self, env):
        self._env = env

    @property
    def net_forces_w(self):
        return self._env.fused.cs_net_forces_w

    @property
    def net_forces_w_history(self):
        return self._env.fused.cs_forces_hist

    @property
    def last_air_time(self):
        return self._env.fused.cs_last_air_time

    @property
    def current_air_time(self):
        return self._env.fused.cs_current_air_time

    @property
    def last_contact_time(self):
        return self._env.fused.cs_last_contact_time

    @property
    def current_contact_time(self):
        return self._env.fused.cs_current_contact_time


class ContactSensorView:
    def __init__(self, env):
        self.data = ContactSensorDataView(env)
        self._env = env

    def compute_first_contact(self, dt: float, abs_tol: float = 1.0e-8):
        c = self.data.current_contact_time   # contact_sensor.py:176-210
        return (c > 0.0) * (c < (dt + abs_tol))

    def compute_first_air(self, dt: float, abs_tol: float = 1.0e-8):
        a = self.data.current_air_time
        return (a > 0.0) * (a < (dt + abs_tol))


class RayCasterDataView:
    def __init__(self, env):
        self._env = env

    @property
    def pos_w(self):
        return self._env.fused.rc_pos_w

    @property
    def quat_w(self):
        return self._env.fused.rc_quat_w

    @property
    def ray_hits_w(self):
        hits = self._env.fused.rc_ray_hits_w
        if hits is None:
            raise AttributeError("ray_hits_w is not materialised (FusedEnv(store_ray_hits=False))")
        return hits


class RayCasterView:
    def __init__(self, env):
        self.data = RayCasterDataView(env)


class SceneView:
    """`env.scene`: `scene["robot"]`, `scene.sensors["contact_forces"]`, `scene.env_origins`, `scene.num_envs`."""

    def __init__(self, env):
        self._env = env
        self.articulations = {"robot": ArticulationView(env)}
        self.sensors = {}
        if env.spec.get("contact"):
            self.sensors["contact_forces"] = ContactSensorView(env)
        if env.spec.get("raycaster"):
            self.sensors["height_scanner"] = RayCasterView(env)

    @property
    def num_envs(self):
        return self._env.num_envs

    @property
    def env_origins(self):
        return self._env.fused.env_origins

    def keys(self):
        return list(self.articulations) + list(self.sensors)

    def __getitem__(self, key):
        if key in self.articulations:
            return self.articulations[key]
        if key in self.sensors:
            return self.sensors[key]
        raise KeyError(f"Scene entity with key '{key}' not found. Available Entities: '{self.keys()}'")
