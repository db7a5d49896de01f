"""The manager objects of `ManagerBasedRLEnv` with the reference's surface (Contract 1) over the fused state.

In the reference each manager loops over its terms in Python every step.  Here the term programs run inside
`b200lab_post_step` / `b200lab_compute_obs`; the managers own no arithmetic.  They expose the same attributes and
methods -- `compute`, `reset(env_ids) -> dict`, `active_terms`, `get_term_cfg` / `set_term_cfg`, `get_term`,
`compute_group`, `get_command` ... -- as live views of the buffers the kernels write (the same tensor object every step,
the reference's ownership convention: reward_manager.py:62,157, termination_manager.py:68-69).

`reset(env_ids)`: the reference resets term state and returns the log entries of the envs being reset
(manager_based_rl_env.py:365-389).  On the fused path the reset itself happened inside the step kernel (masked, no host
sync); `reset` returns the same log entries for the most recent step, read from the device.
"""
from __future__ import annotations

import ctypes

import torch

from .. import _lib


class _Manager:
    def __init__(self, env):
        self._env = env

    @property
    def num_envs(self):
        return self._env.num_envs

    @property
    def device(self):
        return self._env.device

    def _log(self, prefix):
        return {k: v for k, v in self._env.fused.last_step_log().items() if k.startswith(prefix)}


class RewardManager(_Manager):
    """LAB/managers/reward_manager.py."""

    @property
    def active_terms(self):
        return [t["name"] for t in self._env.spec["rewards"]]

    @property
    def _term_names(self):
        return self.active_terms

    @property
    def _episode_sums(self):
        f = self._env.fused
        return {n: f.episode_sums[i] for i, n in enumerate(self.active_terms)}

    @property
    def _step_reward(self):
        return self._env.fused.step_reward[:, :len(self.active_terms)]

    def compute(self, dt=None):
        """The reward buffer of the last step (computed by the step kernel: value * weight * dt, reward_manager.py:128-157)."""
        return self._env.reward_buf

    def reset(self, env_ids=None):
        return self._log("Episode_Reward/")

    def _index(self, term_name):
        if term_name not in self.active_terms:
            raise ValueError(f"Reward term '{term_name}' not found.")   # reward_manager.py:174,188
        return self.active_terms.index(term_name)

    def get_term_cfg(self, term_name):
        """A RewardTermCfg mirroring the live term (reward_manager.py:178-190)."""
        from ..managers import RewardTermCfg
        t = self._env.spec["rewards"][self._index(term_name)]
        src = t.get("cfg")
        return RewardTermCfg(func=src.func if src is not None else t["op"], params=dict(src.params) if src is not None else {},
                             weight=t["weight"])

    def set_term_cfg(self, term_name, cfg):
        """reward_manager.py:163-176.  The weight is re-uploaded to the device descriptor (`b200lab_env_update`); the
        term's function / params are fixed at env construction (they were compiled into the term program)."""
        i = self._index(term_name)
        f = self._env.fused
        self._env.spec["rewards"][i]["weight"] = float(cfg.weight)
        f._st.rew[i].weight = float(cfg.weight)
        _lib.check(f.L.b200lab_env_update(f.handle, ctypes.byref(f._st)))


class TerminationManager(_Manager):
    """LAB/managers/termination_manager.py."""

    @property
    def active_terms(self):
        return [t["name"] for t in self._env.spec["terminations"]]

    @property
    def terminated(self):
        return self._env.reset_terminated

    @property
    def time_outs(self):
        return self._env.reset_time_outs

    @property
    def dones(self):
        return self._env.reset_buf

    def compute(self):
        return self._env.reset_buf

    def get_term(self, name):
        if name not in self.active_terms:
            raise ValueError(f"Termination term '{name}' not found.")
        return self._env.fused.term_dones[self.active_terms.index(name)].view(torch.bool)

    def reset(self, env_ids=None):
        return self._log("Episode_Termination/")


class _VelocityCommandTerm:
    """UniformVelocityCommand's attribute surface (LAB/envs/mdp/commands/velocity_command.py)."""

    def __init__(self, env):
        self._env = env

    @property
    def command(self):
        return self._env.fused.cmd_vel_b

    vel_command_b = command

    @property
    def time_left(self):
        return self._env.fused.cmd_time_left

    @property
    def command_counter(self):
        return self._env.fused.cmd_counter

    @property
    def heading_target(self):
        return self._env.fused.cmd_heading_target

    @property
    def is_heading_env(self):
        return self._env.fused.cmd_is_heading.view(torch.bool)

    @property
    def is_standing_env(self):
        return self._env.fused.cmd_is_standing.view(torch.bool)

    @property
    def metrics(self):
        f = self._env.fused
        return {"error_vel_xy": f.cmd_metric_xy, "error_vel_yaw": f.cmd_metric_yaw}


class CommandManager(_Manager):
    """LAB/managers/command_manager.py."""

    def __init__(self, env):
        super().__init__(env)
        self._term = _VelocityCommandTerm(env) if env.spec.get("command") else None

    @property
    def active_terms(self):
        return [self._env.spec["command"]["name"]] if self._term else []

    def _check(self, name):
        if self._term is None or name != self._env.spec["command"]["name"]:
            raise ValueError(f"Command term '{name}' not found")

    def get_command(self, name):
        self._check(name)
        return self._env.fused.cmd_vel_b

    def get_term(self, name):
        self._check(name)
        return self._term

    def compute(self, dt=None):
        """Metrics, time-left countdown, resampling and the heading controller run inside the step kernel."""

    def reset(self, env_ids=None):
        return self._log("Metrics/")


class _JointActionTerm:
    def __init__(self, env):
        self._env = env

    @property
    def raw_actions(self):
        return self._env.fused.raw_actions

    @property
    def processed_actions(self):
        return self._env.fused.processed_actions

    @property
    def action_dim(self):
        return self._env.fused.A


class ActionManager(_Manager):
    """LAB/managers/action_manager.py."""

    def __init__(self, env):
        super().__init__(env)
        self._term = _JointActionTerm(env)

    @property
    def active_terms(self):
        return [self._env.spec["action"]["name"]]

    @property
    def action(self):
        return self._env.fused.action

    @property
    def prev_action(self):
        return self._env.fused.prev_action

    @property
    def total_action_dim(self):
        return self._env.fused.A

    def get_term(self, name):
        if name != self._env.spec["action"]["name"]:
            raise ValueError(f"Action term '{name}' not found")
        return self._term

    def process_action(self, action):
        self._env.fused.process_action(action)

    def apply_action(self):
        """Folded into `b200lab_substep_pre` (it writes the joint targets before the actuator models read them)."""

    def reset(self, env_ids=None):
        return {}


class ObservationManager(_Manager):
    """LAB/managers/observation_manager.py."""

    @property
    def _group(self):
        return self._env.spec["observations"]["group"]

    @property
    def active_terms(self):
        return {self._group: [t["name"] for t in self._env.spec["observations"]["terms"]]}

    @property
    def group_obs_dim(self):
        return {self._group: (self._env.fused.obs_dim,)}

    @property
    def group_obs_term_dim(self):
        from .term_dims import term_dim
        s = self._env.spec
        return {self._group: [(term_dim(s, t),) for t in s["observations"]["terms"]]}

    @property
    def group_obs_concatenate(self):
        return {self._group: True}

    def compute(self):
        return {self._group: self.compute_group(self._group)}

    def compute_group(self, group_name):
        if group_name != self._group:
            raise ValueError(f"Unable to find the group '{group_name}' in the observation manager."
                             f" Available groups are: {[self._group]}")   # observation_manager.py:277-281
        return self._env.fused.obs_buf

    def reset(self, env_ids=None):
        return {}


class EventManager(_Manager):
    """LAB/managers/event_manager.py.  Reset- and interval-mode terms run inside `b200lab_post_step` (masked per env,
    in the reference's order); startup-mode terms randomise PhysX properties and belong to the simulator."""

    @property
    def available_modes(self):
        ev = self._env.spec.get("events", {})
        modes = []
        if ev.get("reset_root_state") or ev.get("reset_joints"):
            modes.append("reset")
        if ev.get("interval_push"):
            modes.append("interval")
        return modes

    @property
    def active_terms(self):
        ev = self._env.spec.get("events", {})
        out = {}
        if "reset" in self.available_modes:
            out["reset"] = (["reset_root_state_uniform"] if ev.get("reset_root_state") else []) + \
                           [f"reset_joints_by_{e['mode']}" for e in ev.get("reset_joints", [])]
        if ev.get("interval_push"):
            out["interval"] = ["push_by_setting_velocity"] * len(ev["interval_push"])
        return out

    def apply(self, mode, env_ids=None, dt=None, global_env_step_count=None):
        if mode not in self.available_modes:
            return   # event_manager.py:179-182: a mode without terms is a no-op
        raise RuntimeError(f"event mode '{mode}' is applied inside b200lab_post_step, fused with the step it belongs to; "
                           "it cannot be applied on its own from the host")

    def reset(self, env_ids=None):
        return {}


class CurriculumManager(_Manager):
    """LAB/managers/curriculum_manager.py (`terrain_levels_vel` runs inside the step kernel's reset block)."""

    @property
    def active_terms(self):
        c = self._env.spec.get("curriculum")
        return [c["name"]] if c else []

    def compute(self, env_ids=None):
        pass

    def reset(self, env_ids=None):
        return self._log("Curriculum/")
