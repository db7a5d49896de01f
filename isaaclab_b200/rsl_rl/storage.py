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
        self._save_hidden_states(transition.hidden_states)
        self.step += 1

    def _save_hidden_states(self, hidden_states):
        """The recurrent policy's states BEFORE the step, per transition (rollout_storage.py:105-123): lists of
        [T, num_layers, N, H] slabs -- (h, c) for an LSTM, (h,) for a GRU -- for the actor and the critic memory."""
        if hidden_states is None or hidden_states == (None, None):
            return
        T = self.num_transitions_per_env
        per_net = []
        for hs in hidden_states:
            per_net.append(hs if isinstance(hs, (tuple, list)) else (hs,))
        if self.saved_hidden_states_a is None:
            self.saved_hidden_states_a = [torch.zeros(T, *h.shape, device=self.device) for h in per_net[0]]
            self.saved_hidden_states_c = [torch.zeros(T, *h.shape, device=self.device) for h in per_net[1]]
        for slab, h in zip(self.saved_hidden_states_a, per_net[0]):
            slab[self.step].copy_(h)
        for slab, h in zip(self.saved_hidden_states_c, per_net[1]):
            slab[self.step].copy_(h)

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
            # observation rows padded to a multiple of 4 floats (zero pad columns, never written): the first MLP layer reads
            # the [M, D] view of this buffer in place as a 16-byte aligned GEMM operand (modules._Net._input)
            Dp = (D + 3) // 4 * 4
            b = self._mb[M] = dict(obs_block=torch.zeros(M, Dp, device=dev), actions=torch.empty(M, A, device=dev),
                                   values=torch.empty(M, 1, device=dev), returns=torch.empty(M, 1, device=dev),
                                   log_prob=torch.empty(M, 1, device=dev), advantages=torch.empty(M, 1, device=dev),
                                   mu=torch.empty(M, A, device=dev), sigma=torch.empty(M, A, device=dev))
            b["obs"] = b["obs_block"][:, :D]
            b["obs"]._b200_row_padded = Dp != D   # marks the view for the first MLP layer (pad columns stay zero)
        return b

    def gather(self, batch_idx):
        """One fused gather of the eight flattened slabs (rollout_storage.py:188-197)."""
        import ctypes
        M = batch_idx.numel()
        b = self._mb_buffers(M)
        g = _lib.structs["B200LabGather"]()
        g.M, g.obs_dim, g.act_dim = M, b["obs"].shape[1], b["actions"].shape[1]
        g.o_obs_ld = b["obs_block"].shape[1]
        p = _lib.ptr
        g.idx = p(batch_idx)
        g.obs, g.actions, g.values, g.returns = p(self.observations[self._obs_key]), p(self.actions), p(self.values), p(self.returns)
        g.log_prob, g.advantages, g.mu, g.sigma = p(self.actions_log_prob), p(self.advantages), p(self.mu), p(self.sigma)
        g.o_obs, g.o_actions, g.o_values, g.o_returns = p(b["obs_block"]), p(b["actions"]), p(b["values"]), p(b["returns"])
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


    def recurrent_mini_batch_generator(self, num_mini_batches, num_epochs=8):
        """rollout_storage.py:206-260: minibatches are CONTIGUOUS ENV RANGES (whole rollouts, time-major), the observations
        come as zero-padded trajectories [T, n_traj, D] with their validity masks, and the hidden states are those saved
        at each trajectory's first step.  The per-env slabs are returned as views of the storage (no gather)."""
        if self.training_type != "rl":
            raise ValueError("This function is only available for reinforcement learning training.")
        T, N = self.num_transitions_per_env, self.num_envs
        padded_obs, masks, starts = split_and_pad_trajectories(self.observations, self.dones, return_starts=True)
        mb = N // num_mini_batches
        # trajectories are numbered env-major, so an env range owns a contiguous run of them
        per_env = starts.sum(dim=0)                                   # [N] trajectories that begin in each env
        bounds = torch.zeros(num_mini_batches + 1, dtype=torch.int64)
        bounds[1:] = per_env[:num_mini_batches * mb].view(num_mini_batches, mb).sum(1).cumsum(0).cpu()   # one sync per update
        bounds = bounds.tolist()
        start_env_major = starts.t()                                  # [N, T] True where a trajectory begins

        def first_states(slabs):   # [T, L, N, H] -> [n_traj, L, H], trajectory order
            return [sl.permute(2, 0, 1, 3)[start_env_major] for sl in slabs]

        hid_a_all, hid_c_all = first_states(self.saved_hidden_states_a), first_states(self.saved_hidden_states_c)
        for _ in range(num_epochs):
            for i in range(num_mini_batches):
                lo, hi = i * mb, (i + 1) * mb
                t0, t1 = bounds[i], bounds[i + 1]
                hid_a = [h[t0:t1].transpose(1, 0).contiguous() for h in hid_a_all]
                hid_c = [h[t0:t1].transpose(1, 0).contiguous() for h in hid_c_all]
                hid_a = hid_a[0] if len(hid_a) == 1 else hid_a        # a GRU has one state tensor
                hid_c = hid_c[0] if len(hid_c) == 1 else hid_c
                yield (padded_obs[:, t0:t1], self.actions[:, lo:hi], self.values[:, lo:hi], self.advantages[:, lo:hi],
                       self.returns[:, lo:hi], self.actions_log_prob[:, lo:hi], self.mu[:, lo:hi], self.sigma[:, lo:hi],
                       (hid_a, hid_c), masks[:, t0:t1])


def split_and_pad_trajectories(tensor, dones, return_starts=False):
    """RSL/utils/utils.py:78-130.  Cuts every env's rollout at its done flags, stacks the pieces (env-major order) as
    columns of a zero-padded [T, n_traj, ...] tensor and returns it with the [T, n_traj] validity masks:

        [[a1 a2 a3 a4 | a5 a6],        [[a1 a2 a3 a4], [a5 a6 0 0],
         [b1 b2 | b3 b4 b5 | b6]]  ->   [b1 b2 0 0], [b3 b4 b5 0], [b6 0 0 0]]   (shown transposed)

    `tensor`: [T, N, ...] tensor or a TensorDict of such.  Built with index arithmetic on the device (no per-trajectory
    Python split); the one host read is the trajectory count."""
    T, N = dones.shape[0], dones.shape[1]
    d = dones.reshape(T, N) != 0
    starts = torch.ones(T, N, dtype=torch.bool, device=d.device)
    starts[1:] = d[:-1]                                   # a trajectory begins at t = 0 and after every done
    flat_starts = starts.t().reshape(-1)                  # env-major
    traj_id = torch.cumsum(flat_starts.to(torch.int64), 0) - 1
    n_traj = int(traj_id[-1].item()) + 1
    step_in_env = torch.arange(T, device=d.device).repeat(N)
    start_step = step_in_env[flat_starts]                 # [n_traj]
    pos = step_in_env - start_step[traj_id]               # step index inside its trajectory
    lengths = torch.bincount(traj_id, minlength=n_traj)
    masks = lengths.unsqueeze(0) > torch.arange(T, device=d.device).unsqueeze(1)

    def pad(v):
        out = torch.zeros(T, n_traj, *v.shape[2:], device=v.device, dtype=v.dtype)
        out[pos, traj_id] = v.transpose(1, 0).flatten(0, 1)
        return out

    if isinstance(tensor, TensorDict):
        padded = TensorDict({k: pad(v) for k, v in tensor.items()}, batch_size=[T, n_traj], device=dones.device)
    else:
        padded = pad(tensor)
    return (padded, masks, starts) if return_starts else (padded, masks)


def _scratch_doubles(T, N):
    return int(_lib.lib().b200lab_gae_scratch_bytes(T, N) // 8) if _lib.available() else 2 * ((N + 63) // 64) + 2
