"""`PPO` with the rsl_rl 3.0.0 surface (RSL/algorithms/ppo.py:19-469) driving libb200lab kernels.

Per minibatch the reference issues two autograd MLP passes, ~100 eager elementwise kernels, 4 host syncs and
(multi-GPU) three collectives.  Here: gather -> actor/critic forward GEMMs -> ONE fused loss fwd+bwd kernel ->
backward GEMMs into the flat gradient buffer -> ONE all-reduce of [grads | loss stats incl. KL] -> KL-adaptive LR
computed identically on every rank on the device (replaces ppo.py:287-290's broadcast + .item()) -> grad-norm +
clip + Adam in two passes over the flat buffer.  The host reads statistics back once per update().
"""
from __future__ import annotations

import ctypes

import torch

from .. import _lib
from .storage import RolloutStorage


class PPO:
    def __init__(self, policy, num_learning_epochs=5, num_mini_batches=4, clip_param=0.2, gamma=0.99, lam=0.95,
                 value_loss_coef=1.0, entropy_coef=0.01, learning_rate=0.001, max_grad_norm=1.0,
                 use_clipped_value_loss=True, schedule="adaptive", desired_kl=0.01, device="cuda",
                 normalize_advantage_per_mini_batch=False, rnd_cfg=None, symmetry_cfg=None, multi_gpu_cfg=None):
        if rnd_cfg is not None or symmetry_cfg is not None:
            raise NotImplementedError("RND / symmetry augmentation are outside the accelerated path")
        if normalize_advantage_per_mini_batch:
            raise NotImplementedError("normalize_advantage_per_mini_batch=False is the supported configuration")
        self.device = device
        self.is_multi_gpu = multi_gpu_cfg is not None
        self.gpu_global_rank = multi_gpu_cfg["global_rank"] if self.is_multi_gpu else 0
        self.gpu_world_size = multi_gpu_cfg["world_size"] if self.is_multi_gpu else 1
        self.rnd = None
        self.rnd_optimizer = None
        self.symmetry = None
        self.policy = policy
        self.storage: RolloutStorage = None  # type: ignore
        self.transition = RolloutStorage.Transition()
        self.clip_param = clip_param
        self.num_learning_epochs = num_learning_epochs
        self.num_mini_batches = num_mini_batches
        self.value_loss_coef = value_loss_coef
        self.entropy_coef = entropy_coef
        self.gamma = gamma
        self.lam = lam
        self.max_grad_norm = max_grad_norm
        self.use_clipped_value_loss = use_clipped_value_loss
        self.desired_kl = desired_kl
        self.schedule = schedule
        self.normalize_advantage_per_mini_batch = False
        dev = policy.flat_param.device
        self._lr = torch.tensor([learning_rate], device=dev, dtype=torch.float32)
        self.exp_avg = torch.zeros_like(policy.flat_param)
        self.exp_avg_sq = torch.zeros_like(policy.flat_param)
        self.adam_step = 0
        self._sumsq = torch.zeros(1, device=dev)
        self._stat_sum = torch.zeros(8, device=dev)
        self._d_mu = {}
        self.optimizer = _OptimizerView(self)

    # learning rate lives on the device; reading it is a sync (only done for logging)
    @property
    def learning_rate(self):
        return float(self._lr.item())

    @learning_rate.setter
    def learning_rate(self, v):
        self._lr.fill_(float(v))

    def init_storage(self, training_type, num_envs, num_transitions_per_env, obs, actions_shape):
        self.storage = RolloutStorage(training_type, num_envs, num_transitions_per_env, obs, actions_shape, self.device)

    def act(self, obs, eps=None):
        if self.policy.is_recurrent:   # the states the step STARTS from (ppo.py:130-131); the memories replace them in act
            self.transition.hidden_states = self.policy.get_hidden_states()
        self.transition.actions = self.policy.act(obs, eps=eps)
        self.transition.values = self.policy._evaluate(obs)
        self.transition.actions_log_prob = self.policy.get_actions_log_prob(self.transition.actions)
        self.transition.action_mean = self.policy.action_mean
        self.transition.action_sigma = self.policy.action_std
        # The reference keeps `obs` by reference until process_env_step (ppo.py:139) -- safe there because its
        # ObservationManager returns a fresh tensor every step.  The fused env hands back the SAME observation buffer
        # and overwrites it in env.step, so the row of step t is stored now, before the step.
        self.transition.observations = obs
        self.transition.observations_stored = self.storage.store_observations(obs) if self.storage is not None else False
        return self.transition.actions

    def process_env_step(self, obs, rewards, dones, extras):
        self.policy.update_normalization(obs)
        self.transition.rewards = rewards
        self.transition.dones = dones
        self.transition.time_outs = extras.get("time_outs") if extras else None
        # the bootstrap r += gamma * V * time_out (ppo.py:161-164) is fused into the storage copy
        self.storage.add_transitions(self.transition, gamma=self.gamma)
        self.transition.clear()
        self.policy.reset(dones)

    def compute_returns(self, obs):
        last_values = self.policy._evaluate(obs)
        self.storage.compute_returns(last_values, self.gamma, self.lam,
                                     normalize_advantage=not self.normalize_advantage_per_mini_batch)

    def update(self, indices=None):
        pol, L, s = self.policy, _lib.lib(), _lib.current_stream()
        p = _lib.ptr
        n = pol.num_params
        self._stat_sum.zero_()
        num_updates = 0
        inv_world = 1.0 / self.gpu_world_size
        recurrent = pol.is_recurrent
        if recurrent:   # ppo.py:194-197
            generator = self.storage.recurrent_mini_batch_generator(self.num_mini_batches, self.num_learning_epochs)
        else:
            generator = self.storage.mini_batch_generator(self.num_mini_batches, self.num_learning_epochs, indices=indices)
        for (obs_batch, actions_batch, target_values_batch, advantages_batch, returns_batch, old_lp_batch, old_mu_batch,
             old_sigma_batch, hid_states_batch, masks_batch) in generator:
            if recurrent:
                # padded trajectories through the memories from the saved hidden states (ppo.py:249-252), then the
                # [T, n_envs, .] slabs flattened time-major to the rows the MLP / loss kernels see
                x = pol.actor_features(obs_batch, masks_batch, hid_states_batch[0]).contiguous()
                xc = pol.critic_features(obs_batch, masks_batch, hid_states_batch[1]).contiguous()
                flat = lambda v: v.reshape(-1, v.shape[-1]).contiguous()   # noqa: E731
                actions_batch, target_values_batch, advantages_batch = flat(actions_batch), flat(target_values_batch), flat(advantages_batch)
                returns_batch, old_lp_batch = flat(returns_batch), flat(old_lp_batch)
                old_mu_batch, old_sigma_batch = flat(old_mu_batch), flat(old_sigma_batch)
            else:
                x = pol.actor_input(obs_batch)
                xc = pol.critic_input(obs_batch)
            M, A = actions_batch.shape
            mu = pol._a.forward(x, pol.precision)
            value = pol._c.forward(xc, pol.precision)
            d = self._d_mu.get(M)
            if d is None:
                d = self._d_mu[M] = (torch.empty(M, A, device=mu.device), torch.empty(M, 1, device=mu.device))
            pol.flat_grad_ext.zero_()
            a = _lib.structs["B200LabPpoLoss"]()
            a.M, a.A = M, A
            a.mu, a.std, a.value = p(mu), p(pol.std), p(value)
            a.actions, a.old_log_prob, a.old_mu, a.old_sigma = p(actions_batch), p(old_lp_batch), p(old_mu_batch), p(old_sigma_batch)
            a.advantages, a.returns, a.target_values = p(advantages_batch), p(returns_batch), p(target_values_batch)
            a.clip_param, a.value_loss_coef, a.entropy_coef = self.clip_param, self.value_loss_coef, self.entropy_coef
            a.use_clipped_value_loss = 1 if self.use_clipped_value_loss else 0
            a.d_mu, a.d_value = p(d[0]), p(d[1])
            a.d_std = pol.std.grad.data_ptr()
            a.stats = pol.stats.data_ptr()
            _lib.check(L.b200lab_ppo_loss(ctypes.byref(a), s))
            dxa = pol._a.backward(x, d[0], pol.precision, need_dx=recurrent)
            dxc = pol._c.backward(xc, d[1], pol.precision, need_dx=recurrent)
            if recurrent:   # back-propagation through time into memory_a / memory_c
                T = self.storage.num_transitions_per_env
                pol.memory_a.backward(dxa.view(T, -1, dxa.shape[-1]))
                pol.memory_c.backward(dxc.view(T, -1, dxc.shape[-1]))
            if self.is_multi_gpu:
                self.reduce_parameters()
            if self.desired_kl is not None and self.schedule == "adaptive":
                _lib.check(L.b200lab_adapt_lr(p(self._lr), p(pol.stats), float(self.desired_kl), inv_world, s))
            self._sumsq.zero_()
            _lib.check(L.b200lab_grad_sumsq(p(pol.flat_grad), n, p(self._sumsq), s))
            self.adam_step += 1
            _lib.check(L.b200lab_adam_clip_step(p(pol.flat_param), p(pol.flat_grad), p(self.exp_avg), p(self.exp_avg_sq),
                                                n, p(self._lr), p(self._sumsq), inv_world,
                                                float(self.max_grad_norm) if self.max_grad_norm else 0.0,
                                                0.9, 0.999, 1e-8, self.adam_step, s))
            self._stat_sum.add_(pol.stats)
            num_updates += 1
        stats = (self._stat_sum * (inv_world / max(num_updates, 1))).tolist()  # the one host sync of update()
        self.storage.clear()
        return {"value_function": stats[1], "surrogate": stats[0], "entropy": stats[2]}

    # -- multi-GPU (ppo.py:428-469) ------------------------------------------------------------------
    def broadcast_parameters(self):
        torch.distributed.broadcast(self.policy.flat_param, src=0)

    def reduce_parameters(self):
        """ONE all-reduce (SUM) of the flat [grads | stats] buffer; the 1/world_size scaling is folded into the
        grad-norm/Adam kernel and into adapt_lr."""
        torch.distributed.all_reduce(self.policy.flat_grad_ext, op=torch.distributed.ReduceOp.SUM)


class _OptimizerView:
    """Checkpoint-compatible facade: `state_dict()/load_state_dict()` in torch.optim.Adam's format so
    OnPolicyRunner.save/load interchange with the reference (on_policy_runner.py:289-324)."""

    def __init__(self, alg: PPO):
        self.alg = alg

    @property
    def param_groups(self):
        return [{"lr": self.alg.learning_rate}]

    def state_dict(self):
        alg, pol = self.alg, self.alg.policy
        state = {}
        params = list(pol.parameters())

        def as_param(buf, i, off, k):   # the slice of a flat buffer in parameter i's shape (K-padded blocks: drop the pad)
            shape = params[i].shape
            if k != params[i].numel():
                return buf[off:off + k].view(shape[0], k // shape[0])[:, :shape[1]]
            return buf[off:off + k].view(shape)

        for i, (off, k) in enumerate(pol._views):
            state[i] = {"step": torch.tensor(float(alg.adam_step)),
                        "exp_avg": as_param(alg.exp_avg, i, off, k).clone(),
                        "exp_avg_sq": as_param(alg.exp_avg_sq, i, off, k).clone()}
        group = {"lr": alg.learning_rate, "betas": (0.9, 0.999), "eps": 1e-8, "weight_decay": 0, "amsgrad": False,
                 "params": list(range(len(pol._views)))}
        return {"state": state, "param_groups": [group]}

    def load_state_dict(self, sd):
        alg, pol = self.alg, self.alg.policy
        params = list(pol.parameters())
        for i, (off, k) in enumerate(pol._views):
            st = sd["state"].get(i)
            if st is None:
                continue
            shape = params[i].shape
            for buf, key in ((alg.exp_avg, "exp_avg"), (alg.exp_avg_sq, "exp_avg_sq")):
                if k != params[i].numel():   # K-padded block
                    buf[off:off + k].view(shape[0], k // shape[0])[:, :shape[1]].copy_(st[key].reshape(shape))
                else:
                    buf[off:off + k].copy_(st[key].reshape(-1))
            alg.adam_step = int(float(st["step"]))
        if sd.get("param_groups"):
            alg.learning_rate = sd["param_groups"][0]["lr"]
