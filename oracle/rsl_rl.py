"""Oracle restatement of the rsl_rl 3.0.0 hot path in plain torch on CPU (test infrastructure only).

RSL = /root/reference/source/rsl_rl/rsl_rl
"""
from __future__ import annotations

import torch


def compute_returns(rewards, values, dones, last_values, gamma, lam, normalize_advantage=True):
    """RSL/storage/rollout_storage.py:127-149.  rewards/values [T,N,1] f32, dones [T,N,1] u8, last_values [N,1]."""
    T = rewards.shape[0]
    returns = torch.zeros_like(rewards)
    advantage = 0
    for step in reversed(range(T)):
        next_values = last_values if step == T - 1 else values[step + 1]
        next_is_not_terminal = 1.0 - dones[step].float()
        delta = rewards[step] + next_is_not_terminal * gamma * next_values - values[step]
        advantage = delta + next_is_not_terminal * gamma * lam * advantage
        returns[step] = advantage + values[step]
    advantages = returns - values
    if normalize_advantage:
        advantages = (advantages - advantages.mean()) / (advantages.std() + 1e-8)
    return returns, advantages


def bootstrap_rewards(rewards, values, time_outs, gamma):
    """RSL/algorithms/ppo.py:161-164.  rewards [N], values [N,1], time_outs [N] (bool/float)."""
    return rewards + gamma * torch.squeeze(values * time_outs.unsqueeze(1).to(values.dtype), 1)


def mini_batch_slices(indices, num_mini_batches, num_epochs, batch_size):
    """RSL/storage/rollout_storage.py:160-203: the same permutation is reused for every epoch."""
    mb = batch_size // num_mini_batches
    for _ in range(num_epochs):
        for i in range(num_mini_batches):
            yield indices[i * mb:(i + 1) * mb]


def mlp_forward(params, prefix, x, n_layers):
    """RSL/networks/mlp.py:60-80: Linear (+ELU) ... Linear; state-dict names `<prefix>.{0,2,4,...}.{weight,bias}`."""
    for l in range(n_layers):
        x = torch.nn.functional.linear(x, params[f"{prefix}.{2 * l}.weight"], params[f"{prefix}.{2 * l}.bias"])
        if l < n_layers - 1:
            x = torch.nn.functional.elu(x)
    return x


def n_layers_of(params, prefix):
    return len([k for k in params if k.startswith(prefix + ".") and k.endswith(".weight")])


def normal_log_prob(actions, mu, sigma):
    """torch.distributions.Normal.log_prob summed over the action dim (RSL/modules/actor_critic.py:143)."""
    var = sigma ** 2
    return (-((actions - mu) ** 2) / (2 * var) - sigma.log() - 0.5 * torch.log(torch.tensor(2 * torch.pi))).sum(-1)


def normal_entropy(sigma):
    """Normal.entropy().sum(-1) (actor_critic.py:99)."""
    return (0.5 + 0.5 * torch.log(torch.tensor(2 * torch.pi)) + sigma.log()).sum(-1)


def act(params, obs, eps):
    """PPO.act (ppo.py:129-140) with the standard-normal draws `eps` made explicit: a = mu + std * eps."""
    mu = mlp_forward(params, "actor", obs, n_layers_of(params, "actor"))
    sigma = params["std"].expand_as(mu)
    actions = mu + sigma * eps
    values = mlp_forward(params, "critic", obs, n_layers_of(params, "critic"))
    return actions, values, normal_log_prob(actions, mu, sigma), mu, sigma


def ppo_loss(params, batch, clip_param, value_loss_coef, entropy_coef, use_clipped_value_loss=True):
    """ppo.py:249-315 (no symmetry / RND).  batch: obs, actions, target_values, advantages, returns,
    old_log_prob, old_mu, old_sigma (shapes as yielded by mini_batch_generator)."""
    mu = mlp_forward(params, "actor", batch["obs"], n_layers_of(params, "actor"))
    sigma = params["std"].expand_as(mu)
    log_prob = normal_log_prob(batch["actions"], mu, sigma)
    value = mlp_forward(params, "critic", batch["obs"], n_layers_of(params, "critic"))
    entropy = normal_entropy(sigma)
    with torch.no_grad():
        kl = torch.sum(torch.log(sigma / batch["old_sigma"] + 1.0e-5)
                       + (torch.square(batch["old_sigma"]) + torch.square(batch["old_mu"] - mu))
                       / (2.0 * torch.square(sigma)) - 0.5, axis=-1)
        kl_mean = torch.mean(kl)
    ratio = torch.exp(log_prob - torch.squeeze(batch["old_log_prob"]))
    adv = torch.squeeze(batch["advantages"])
    surrogate = -adv * ratio
    surrogate_clipped = -adv * torch.clamp(ratio, 1.0 - clip_param, 1.0 + clip_param)
    surrogate_loss = torch.max(surrogate, surrogate_clipped).mean()
    if use_clipped_value_loss:
        value_clipped = batch["target_values"] + (value - batch["target_values"]).clamp(-clip_param, clip_param)
        value_loss = torch.max((value - batch["returns"]).pow(2), (value_clipped - batch["returns"]).pow(2)).mean()
    else:
        value_loss = (batch["returns"] - value).pow(2).mean()
    loss = surrogate_loss + value_loss_coef * value_loss - entropy_coef * entropy.mean()
    return loss, dict(surrogate=surrogate_loss.detach(), value_function=value_loss.detach(),
                      entropy=entropy.mean().detach(), kl=kl_mean, mu=mu, value=value)


def adapt_lr(lr, kl_mean, desired_kl):
    """ppo.py:281-284."""
    if kl_mean > desired_kl * 2.0:
        return max(1e-5, lr / 1.5)
    if kl_mean < desired_kl / 2.0 and kl_mean > 0.0:
        return min(1e-2, lr * 1.5)
    return lr


class Adam:
    """torch.optim.Adam defaults (betas 0.9/0.999, eps 1e-8, no weight decay / amsgrad), restated."""

    def __init__(self, params):
        self.m = {k: torch.zeros_like(v) for k, v in params.items()}
        self.v = {k: torch.zeros_like(v) for k, v in params.items()}
        self.t = 0

    def step(self, params, grads, lr, b1=0.9, b2=0.999, eps=1e-8):
        self.t += 1
        bc1, bc2 = 1 - b1 ** self.t, 1 - b2 ** self.t
        for k in params:
            g = grads[k]
            self.m[k].lerp_(g, 1 - b1)
            self.v[k].mul_(b2).addcmul_(g, g, value=1 - b2)
            denom = (self.v[k].sqrt() / (bc2 ** 0.5)).add_(eps)
            params[k].addcdiv_(self.m[k], denom, value=-lr / bc1)


def clip_grad_norm(grads, max_norm):
    """torch.nn.utils.clip_grad_norm_ (L2)."""
    total = torch.linalg.vector_norm(torch.stack([torch.linalg.vector_norm(g) for g in grads.values()]))
    coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
    for g in grads.values():
        g.mul_(coef)
    return total


def ppo_update(params, storage, indices, cfg, lr, world_size=1):
    """PPO.update (ppo.py:178-422) on flattened [T*N,.] storage tensors with an explicit permutation.
    `params` (dict of leaf tensors) is updated in place.  Returns (loss_dict, lr, lr_per_update, adam)."""
    flat = {k: v.flatten(0, 1) for k, v in storage.items()}
    B = flat["actions"].shape[0]
    adam = Adam(params)
    sums = dict(value_function=0.0, surrogate=0.0, entropy=0.0)
    lrs = []
    n_upd = 0
    for idx in mini_batch_slices(indices, cfg["mini_batches"], cfg["epochs"], B):
        batch = dict(obs=flat["observations"][idx], actions=flat["actions"][idx], target_values=flat["values"][idx],
                     advantages=flat["advantages"][idx], returns=flat["returns"][idx],
                     old_log_prob=flat["actions_log_prob"][idx], old_mu=flat["mu"][idx], old_sigma=flat["sigma"][idx])
        leaves = {k: v.detach().requires_grad_(True) for k, v in params.items()}
        loss, info = ppo_loss(leaves, batch, cfg["clip_param"], cfg["value_loss_coef"], cfg["entropy_coef"])
        if cfg.get("desired_kl") is not None:
            lr = adapt_lr(lr, info["kl"], cfg["desired_kl"])
        loss.backward()
        grads = {k: v.grad for k, v in leaves.items()}
        clip_grad_norm(grads, cfg["max_grad_norm"])
        lrs.append(lr)
        with torch.no_grad():
            adam.step(params, grads, lr)
        for k in sums:
            sums[k] += float(info[k])
        n_upd += 1
    return {k: v / n_upd for k, v in sums.items()}, lr, lrs, adam


class EmpiricalNormalization:
    """RSL/networks/normalization.py:14-72 restated: running mean / biased variance over all samples seen."""

    def __init__(self, dim, eps=1e-2, until=None):
        self.eps, self.until = eps, until
        self.mean, self.var, self.std = torch.zeros(1, dim), torch.ones(1, dim), torch.ones(1, dim)
        self.count = 0

    def __call__(self, x):
        return (x - self.mean) / (self.std + self.eps)          # normalization.py:47

    def update(self, x):
        if self.until is not None and self.count >= self.until:  # normalization.py:55-56
            return
        n = x.shape[0]
        self.count += n
        rate = n / self.count
        var_x = torch.var(x, dim=0, unbiased=False, keepdim=True)
        mean_x = torch.mean(x, dim=0, keepdim=True)
        delta = mean_x - self.mean
        self.mean = self.mean + rate * delta                      # normalization.py:64
        self.var = self.var + rate * (var_x - self.var + delta * (mean_x - self.mean))   # :65 (uses the NEW mean)
        self.std = torch.sqrt(self.var)
