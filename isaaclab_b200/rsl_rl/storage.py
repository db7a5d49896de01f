"""`RolloutStorage` with the rsl_rl 3.0.0 surface (RSL/storage/rollout_storage.py:13-203) on libb200lab kernels.

Buffers keep the reference layout `[T, N, .]` fp32 / uint8 in HBM.  `add_transitions` is one fused copy kernel
(including PPO's time-out bootstrap when `time_outs` is passed), `compute_returns` one scan kernel + one
normalisation kernel, and each minibatch of `mini_batch_generator` one gather kernel over the eight slabs.
"""
from __future__ import annotations

import torch

from .. import _lib
from .tensordict import TensorDict


class RolloutStorage:
    class Transition:
        def __init__(self):
            self.observations = None
            self.actions = None
            self.privileged_actions = None
            self.rewards = None
            self.dones = None
            self.values = None
            self.actions_log_prob = None
            self.action_mean = None
            self.action_sigma = None
            self.hidden_states = None
            self.time_outs = None  # b200 extension: fused bootstrap (ppo.py:161-164)
            self.observations_stored = False  # the observation row of this step is already in the slab (PPO.act)

        def clear(self):
            self.__init__()

    def __init__(self, training_type, num_envs, num_transitions_per_env, obs, actions_shape, device="cuda"):
        if training_type != "rl":
            raise ValueError("isaaclab_b200 RolloutStorage implements training_type='rl'")
        self.training_type = training_type
        self.device = device
        self.num_transitions_per_env = T = num_transitions_per_env
        self.num_envs = N = num_envs
        self.actions_shape = actions_shape
        keys = list(obs.keys())
        # several observation groups: one slab per group; the fused gather kernel moves the first one, the others are
        # gathered per minibatch with one indexed copy each (mini_batch_generator)
        self._obs_key = keys[0]
        self.observations = TensorDict({k: torch.zeros(T, *obs[k].shape, device=device) for k in keys},
                                       batch_size=[T, N], device=device)
        A = int(actions_shape[0])
        self.rewards = torch.zeros(T, N, 1, device=device)
        self.actions = torch.zeros(T, N, A, device=device)
        self.dones = torch.zeros(T, N, 1, device=device, dtype=torch.uint8)
        self.values = torch.zeros(T, N, 1, device=device)
        self.actions_log_prob = torch.zeros(T, N, 1, device=device)
        self.mu = torch.zeros(T, N, A, device=device)
        self.sigma = torch.zeros(T, N, A, device=device)
        self.returns = torch.zeros(T, N, 1, device=device)
        self.advantages = torch.zeros(T, N, 1, device=device)
        self.saved_hidden_states_a = None
        self.saved_hidden_states_c = None
        self.step = 0
        self._gae_scratch = torch.zeros(max(1, _scratch_doubles(T, N)), dtype=torch.float64, device=device)
        self._mb = {}

    def store_observations(self, obs) -> bool:
        """Copies the observation of the CURRENT step into the slab (every group).  Called by `PPO.act`: the fused env
        returns one persistent observation buffer that `env.step` overwrites, so the row has to be taken before the
        step, not in `add_transitions` after it (rollout_storage.py:86 copies a tensor the reference env never reuses)."""
        if self.step >= self.num_transitions_per_env:
            raise OverflowError("Rollout buffer overflow! You should call clear() before adding new transitions.")
        for k in self.observations.keys():
            self.observations[k][self.step].copy_(obs[k])
        return True

    def add_transitions(self, transition: "RolloutStorage.Transition", gamma: float = 0.0):
        if self.step >= self.num_transitions_per_env:
            raise OverflowError("Rollout buffer overflow! You should call clear() before adding new transitions.")
        S = _lib.structs["B200LabTransition"]
        obs = transition.observations[self._obs_key]
        tr = S()
        tr.T, tr.N = self.num_transitions_per_env, self.num_envs
        tr.obs_dim, tr.act_dim, tr.t = obs.shape[-1], self.actions.shape[-1], self.step
        tr.gamma = gamma
        dones = transition.dones
        if dones.dtype != torch.uint8:
            dones = dones.to(torch.uint8)
        touts = transition.time_outs
        if touts is not None and touts.dtype != torch.uint8:
            touts = touts.to(torch.uint8)
        sigma = transition.action_sigma
        if not sigma.is_contiguous():
            sigma = sigma.contiguous()
        p = _lib.ptr
        # locals keep every (possibly temporary) tensor alive until the kernel is enqueued: the C ABI borrows pointers
        obs_c, rew_c, dones_c = obs.contiguous(), transition.rewards.contiguous(), dones.contiguous()
        touts_c = touts.contiguous() if touts is not None else None
        lp_c = transition.actions_log_prob.contiguous()
        stored = bool(getattr(transition, "observations_stored", False))
        tr.obs, tr.actions, tr.rewards = (None if stored else p(obs_c)), p(transition.actions), p(rew_c)
        tr.dones, tr.time_outs, tr.values = p(dones_c), p(touts_c), p(transition.values)
        tr.log_prob, tr.mu, tr.sigma = p(lp_c), p(transition.action_mean), p(sigma)
        tr.s_obs, tr.s_actions, tr.s_rewards = p(self.observations[self._obs_key]), p(self.actions), p(self.rewards)
        tr.s_dones, tr.s_values, tr.s_log_prob = p(self.dones), p(self.values), p(self.actions_log_prob)
        tr.s_mu, tr.s_sigma = p(self.mu), p(self.sigma)
        import ctypes
        _lib.check(_lib.lib().b200lab_add_transition(ctypes.byref(tr), _lib.current_stream()))
        if not stored:
            for k in self.observations.keys():
                if k != self._obs_key:
                    self.observations[k][self.step].copy_(transition.observations[k])
        self.step += 1

    def clear(self):
        self.step = 0

    def compute_returns(self, last_values, gamma, lam, normalize_advantage: bool = True):
        T, N = self.num_transitions_per_env, self.num_envs
        p = _lib.ptr
        last_values = last_values.contiguous()
        _lib.check(_lib.lib().b200lab_gae(p(self.rewards), p(self.values), p(self.dones), p(last_values),
                                          p(self.returns), p(self.advantages), T, N, float(gamma), float(lam),
                                          1 if normalize_advantage else 0, p(self._gae_scratch),
                                          _lib.current_stream()))

    def _mb_buffers(self, M):
        b = self._mb.get(M)
        if b is None:
            D, A, dev = self.observations[self._obs_key].shape[-1], self.actions.shape[-1], self.device
            b = self._mb[M] = dict(obs=torch.empty(M, D, device=dev), actions=torch.empty(M, A, device=dev),
                                   values=torch.empty(M, 1, device=dev), returns=torch.empty(M, 1, device=dev),
                                   log_prob=torch.empty(M, 1, device=dev), advantages=torch.empty(M, 1, device=dev),
                                   mu=torch.empty(M, A, device=dev), sigma=torch.empty(M, A, device=dev))
        return b

    def gather(self, batch_idx):
        """One fused gather of the eight flattened slabs (rollout_storage.py:188-197)."""
        import ctypes
        M = batch_idx.numel()
        b = self._mb_buffers(M)
        g = _lib.structs["B200LabGather"]()
        g.M, g.obs_dim, g.act_dim = M, b["obs"].shape[1], b["actions"].shape[1]
        p = _lib.ptr
        g.idx = p(batch_idx)
        g.obs, g.actions, g.values, g.returns = p(self.observations[self._obs_key]), p(self.actions), p(self.values), p(self.returns)
        g.log_prob, g.advantages, g.mu, g.sigma = p(self.actions_log_prob), p(self.advantages), p(self.mu), p(self.sigma)
        g.o_obs, g.o_actions, g.o_values, g.o_returns = p(b["obs"]), p(b["actions"]), p(b["values"]), p(b["returns"])
        g.o_log_prob, g.o_advantages, g.o_mu, g.o_sigma = p(b["log_prob"]), p(b["advantages"]), p(b["mu"]), p(b["sigma"])
        _lib.check(_lib.lib().b200lab_gather_minibatch(ctypes.byref(g), _lib.current_stream()))
        return b

    def mini_batch_generator(self, num_mini_batches, num_epochs=8, indices=None):
        """Yields the reference's 10-tuple.  `indices` (int64 permutation) may be supplied for parity runs;
        otherwise `torch.randperm` on the device, drawn once and reused by every epoch (rollout_storage.py:165)."""
        batch_size = self.num_envs * self.num_transitions_per_env
        mini_batch_size = batch_size // num_mini_batches
        if indices is None:
            indices = torch.randperm(num_mini_batches * mini_batch_size, device=self.device)
        indices = indices.to(device=self.device, dtype=torch.int64).contiguous()
        for _ in range(num_epochs):
            for i in range(num_mini_batches):
                idx = indices[i * mini_batch_size:(i + 1) * mini_batch_size]
                b = self.gather(idx)
                groups = {self._obs_key: b["obs"]}
                for k in self.observations.keys():   # further observation groups (e.g. an asymmetric critic's set)
                    if k != self._obs_key:
                        groups[k] = self.observations[k].flatten(0, 1)[idx]
                obs_batch = TensorDict(groups, batch_size=[mini_batch_size], device=self.device)
                yield (obs_batch, b["actions"], b["values"], b["advantages"], b["returns"], b["log_prob"], b["mu"],
                       b["sigma"], (None, None), None)


def _scratch_doubles(T, N):
    return int(_lib.lib().b200lab_gae_scratch_bytes(T, N) // 8) if _lib.available() else 2 * ((N + 63) // 64) + 2
