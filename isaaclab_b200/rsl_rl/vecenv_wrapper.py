"""`RslRlVecEnvWrapper`: the boundary adapter between `ManagerBasedRLEnv` and the rsl_rl runner
(LABRL/rsl_rl/vecenv_wrapper.py:16-195 for the attribute surface, RSL/env/vec_env.py:13-98 for the 3.0.0 contract the
vendored runner actually consumes).

The reference tree pairs a 2.3.1-era wrapper (`get_observations() -> (tensor, extras)`) with the 3.0.0 runner
(`get_observations() -> TensorDict`); SURVEY.md notes the mismatch.  This wrapper serves the 3.0.0 contract and
offers the tuple form through `legacy_api=True`.

Nothing is copied per step: the observation TensorDict wraps the env's persistent observation buffers and the dones
tensor is a persistent buffer refreshed by one copy kernel.
"""
from __future__ import annotations

import torch

from .tensordict import TensorDict


class RslRlVecEnvWrapper:
    def __init__(self, env, clip_actions: float | None = None, legacy_api: bool = False):
        base = getattr(env, "unwrapped", env)
        for attr in ("num_envs", "max_episode_length", "step", "reset", "episode_length_buf"):
            if not hasattr(base, attr):
                raise ValueError("The environment must provide the ManagerBasedRLEnv interface. Environment type: "
                                 f"{type(env)}")
        self.env = env
        self.clip_actions = clip_actions
        self.legacy_api = legacy_api
        self.num_envs = base.num_envs
        self.device = base.device
        self.max_episode_length = base.max_episode_length
        self.num_actions = base.action_manager.total_action_dim
        dims = base.observation_manager.group_obs_dim
        self.num_obs = dims["policy"][0]
        self.num_privileged_obs = dims["critic"][0] if "critic" in dims else 0
        self._dones = torch.zeros(self.num_envs, dtype=torch.long, device=self.device)
        self._clipped = None
        self._obs_td = None
        # the runner never calls reset (vecenv_wrapper.py:77-78)
        self.env.reset()

    def __str__(self):
        return f"<{type(self).__name__}{self.env}>"

    __repr__ = __str__

    @classmethod
    def class_name(cls) -> str:
        return cls.__name__

    @property
    def unwrapped(self):
        return getattr(self.env, "unwrapped", self.env)

    @property
    def cfg(self):
        return self.unwrapped.cfg

    @property
    def episode_length_buf(self) -> torch.Tensor:
        return self.unwrapped.episode_length_buf

    @episode_length_buf.setter
    def episode_length_buf(self, value: torch.Tensor):
        self.unwrapped.episode_length_buf = value

    # -- observations -----------------------------------------------------------------------------
    def _as_td(self, obs_dict):
        td = self._obs_td
        if td is None or any(td[k] is not v for k, v in obs_dict.items()):
            td = self._obs_td = TensorDict(dict(obs_dict), batch_size=[self.num_envs], device=self.device)
        return td

    def get_observations(self):
        obs_dict = self.unwrapped.observation_manager.compute()
        if self.legacy_api:
            return obs_dict["policy"], {"observations": obs_dict}
        return self._as_td(obs_dict)

    # -- MDP ----------------------------------------------------------------------------------------
    def seed(self, seed: int = -1) -> int:
        return self.unwrapped.seed(seed)

    def reset(self):
        obs_dict, _ = self.env.reset()
        if self.legacy_api:
            return obs_dict["policy"], {"observations": obs_dict}
        return self._as_td(obs_dict), {"observations": obs_dict}

    def step(self, actions: torch.Tensor):
        if self.clip_actions is not None:
            if self._clipped is None or self._clipped.shape != actions.shape or self._clipped.device != actions.device:
                self._clipped = torch.empty_like(actions)
            actions = torch.clamp(actions, -self.clip_actions, self.clip_actions, out=self._clipped)
        obs_dict, rew, terminated, truncated, extras = self.env.step(actions)
        base = self.unwrapped
        # dones = (terminated | truncated).long() (vecenv_wrapper.py:178); the env already holds the OR as reset_buf
        reset_buf = getattr(base, "reset_buf", None)
        if reset_buf is not None:
            self._dones.copy_(reset_buf)
        else:
            torch.logical_or(terminated, truncated, out=self._dones)
        extras["observations"] = obs_dict
        if not getattr(base.cfg, "is_finite_horizon", False):
            extras["time_outs"] = truncated
        obs = obs_dict["policy"] if self.legacy_api else self._as_td(obs_dict)
        return obs, rew, self._dones, extras

    def episode_log(self) -> dict:
        """`extras["log"]` aggregates since the last call, read with one device->host copy
        (ManagerBasedRLEnv.episode_log)."""
        base = self.unwrapped
        return base.episode_log() if hasattr(base, "episode_log") else {}

    def close(self):
        return self.env.close() if hasattr(self.env, "close") else None
