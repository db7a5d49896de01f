"""`ActorCritic` with the rsl_rl 3.0.0 interface (RSL/modules/actor_critic.py:15-168) on libb200lab kernels.

Same constructor kwargs, same method surface (`act`, `evaluate`, `act_inference`, `get_actions_log_prob`,
`action_mean`, `action_std`, `entropy`, `update_normalization`, `reset`) and the same `state_dict()` keys
(`actor.{0,2,..}.{weight,bias}`, `critic.*`, `std`) so checkpoints interchange with the reference
(RSL/runners/on_policy_runner.py:289-324).

All parameters are views into ONE flat fp32 buffer (and their `.grad`s into one flat gradient buffer), so the
gradient all-reduce, the grad-norm and Adam are each a single pass over contiguous memory.  There is no autograd
on this path: forward and backward are explicit kernel calls through the C ABI (include/b200lab.h).
"""
from __future__ import annotations

import math

import torch
from torch import nn

from .. import _lib


def _mlp(in_dim, out_dim, hidden):
    layers, d = [], in_dim
    for h in hidden:
        h = in_dim if h == -1 else h
        layers += [nn.Linear(d, h), nn.ELU()]
        d = h
    layers.append(nn.Linear(d, out_dim))
    return nn.Sequential(*layers)  # module names "0","1",... as RSL/networks/mlp.py:77-80


def _padded_k(weight):
    """Row stride of a 2-D weight's storage block: in_features, or the next multiple of 4 when `_flatten` padded it."""
    return int(weight.stride(0)) if weight.dim() == 2 else int(weight.shape[-1])


_BWD_ELU_OUT, _BWD_ELU_IN, _BWD_DZ_GIVEN = 1, 2, 4   # include/b200lab.h: B200LAB_BWD_*


class _Net:
    """Per-network forward/backward driver over raw pointers."""

    def __init__(self, seq: nn.Sequential):
        self.linears = [m for m in seq if isinstance(m, nn.Linear)]
        self.in_features = self.linears[0].in_features
        # the first layer's weight may live in a K-padded block (ActorCritic._flatten): its GEMMs then run on the padded
        # width, against an input padded with zero columns
        self.dims = [_padded_k(self.linears[0].weight)] + [l.out_features for l in self.linears]
        self._cache = {}

    def _input(self, x, b):
        """x [M, in_features] -> the [M, dims[0]] operand of layer 0 (zero columns beyond in_features)."""
        if self.dims[0] == self.in_features:
            return x
        if x.stride(0) == self.dims[0] and x.stride(1) == 1 and x.data_ptr() % 16 == 0 and getattr(x, "_b200_row_padded", False):
            b["x0"] = x          # already the [M, in_features] view of a zero-padded [M, dims[0]] block (RolloutStorage.gather)
            return x
        xp = b.get("xpad")
        if xp is None:
            xp = b["xpad"] = torch.zeros(x.shape[0], self.dims[0], device=x.device)
        xp[:, :self.in_features].copy_(x)
        b["x0"] = xp
        return xp

    def buffers(self, M, device):
        key = (M, str(device))
        b = self._cache.get(key)
        if b is None:
            if len(self._cache) > 4:
                self._cache.clear()
            acts = [torch.empty(M, d, device=device) for d in self.dims[1:]]
            grads = [torch.empty(M, d, device=device) for d in self.dims[1:]]
            ws = max(_lib.lib().b200lab_linear_bwd_workspace_bytes(M, self.dims[i], self.dims[i + 1])
                     for i in range(len(self.linears)))
            work = torch.empty((ws + 3) // 4, device=device)
            b = self._cache[key] = dict(acts=acts, grads=grads, work=work)
        return b

    def forward(self, x, precision):
        L = _lib.lib()
        M = x.shape[0]
        b = self.buffers(M, x.device)
        s = _lib.current_stream()
        inp = self._input(x, b)
        for i, lin in enumerate(self.linears):
            last = i == len(self.linears) - 1
            # weights by raw address: a K-padded first-layer weight is a strided view of its (contiguous) block
            _lib.check(L.b200lab_linear_fwd(inp.data_ptr(), lin.weight.data_ptr(), _lib.ptr(lin.bias),
                                            _lib.ptr(b["acts"][i]), M, self.dims[i], self.dims[i + 1],
                                            0 if last else 1, precision, s))
            inp = b["acts"][i]
        return inp

    def backward(self, x, d_out, precision, need_dx=False):
        """Gradients of the last forward(x); d_out is dLoss/dOutput [M, out]; fills every layer's .grad view.  With
        `need_dx` the gradient w.r.t. `x` is computed too and returned (the recurrent policy feeds it to the RNN)."""
        L = _lib.lib()
        M = x.shape[0]
        b = self.buffers(M, x.device)
        s = _lib.current_stream()
        dy = d_out
        if need_dx and "dx0" not in b:
            b["dx0"] = torch.empty(M, self.dims[0], device=x.device)
        x0 = b["x0"] if self.dims[0] != self.in_features else x   # the padded operand forward(x) used
        # every layer below the last one receives its dZ (and its bias gradient) from the epilogue of the dX product of
        # the layer above it (B200LAB_BWD_ELU_IN there, B200LAB_BWD_DZ_GIVEN here): no stand-alone ELU-backward pass
        for i in reversed(range(len(self.linears))):
            lin = self.linears[i]
            last = i == len(self.linears) - 1
            inp = x0 if i == 0 else b["acts"][i - 1]
            dx = (b["dx0"] if need_dx else None) if i == 0 else b["grads"][i - 1]
            flags = (0 if last else _BWD_DZ_GIVEN) | (_BWD_ELU_IN if i > 0 else 0)
            db_in = self.linears[i - 1].bias.grad if i > 0 else None
            _lib.check(L.b200lab_linear_bwd_chain(inp.data_ptr(), lin.weight.data_ptr(), None, _lib.ptr(dy), None,
                                                  _lib.ptr(dx), lin.weight.grad.data_ptr(), _lib.ptr(lin.bias.grad),
                                                  _lib.ptr(db_in), M, self.dims[i], self.dims[i + 1], flags, precision,
                                                  _lib.ptr(b["work"]), s))
            dy = dx
        return b["dx0"][:, :self.in_features] if need_dx else None


class EmpiricalNormalization(nn.Module):
    """Running mean / variance normalisation of observations (RSL/networks/normalization.py:14-72): same buffers
    (`_mean`, `_var`, `_std` of shape [1, D], `count`) so state dicts interchange; `forward` and `update` are one kernel
    each.  The sample counter is mirrored on the host (the batch size is known there), so neither call synchronises."""

    def __init__(self, shape, eps=1e-2, until=None, device="cuda", global_stats=True):
        """`global_stats`: in a multi-GPU run (torch.distributed initialised, world size > 1) every update learns the
        UNION of the ranks' batches -- one all-reduce of 2D+1 doubles -- so all ranks normalise identically (the
        reference updates rank-locally and lets the statistics drift apart).  No effect with one process."""
        super().__init__()
        self.eps, self.until = eps, until
        self.global_stats = bool(global_stats)
        self._sums = None
        D = int(shape if isinstance(shape, int) else shape[0])
        self.register_buffer("_mean", torch.zeros(1, D, device=device))
        self.register_buffer("_var", torch.ones(1, D, device=device))
        self.register_buffer("_std", torch.ones(1, D, device=device))
        self.register_buffer("count", torch.tensor(0, dtype=torch.long, device=device))
        self._count = 0
        self._out = {}

    @property
    def mean(self):
        return self._mean.squeeze(0).clone()

    @property
    def std(self):
        return self._std.squeeze(0).clone()

    def forward(self, x):
        x = x.contiguous()
        M, D = x.shape
        out = self._out.get(M)
        if out is None:
            if len(self._out) > 4:
                self._out.clear()
            out = self._out[M] = torch.empty(M, D, device=x.device)
        p = _lib.ptr
        _lib.check(_lib.lib().b200lab_empirical_norm_apply(p(x), p(self._mean), p(self._std), float(self.eps), p(out),
                                                          M, D, _lib.current_stream()))
        return out

    def update(self, x):
        if not self.training:
            return
        if self.until is not None and self._count >= self.until:
            return
        x = x.contiguous()
        M, D = x.shape
        p = _lib.ptr
        dist = torch.distributed
        if self.global_stats and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            if self._sums is None:
                self._sums = torch.zeros(2 * D + 1, dtype=torch.float64, device=x.device)
            L, s = _lib.lib(), _lib.current_stream()
            _lib.check(L.b200lab_empirical_norm_partial(p(x), M, D, p(self._sums), s))
            dist.all_reduce(self._sums, op=dist.ReduceOp.SUM)
            _lib.check(L.b200lab_empirical_norm_merge(p(self._sums), D, self._count, p(self._mean), p(self._var), p(self._std), s))
            self._count += M * dist.get_world_size()   # env shards are equal-sized; the kernels use the exact reduced count
            self.count.fill_(self._count)
            return
        self._count += M
        self.count.fill_(self._count)
        _lib.check(_lib.lib().b200lab_empirical_norm_update(p(x), M, D, self._count, p(self._mean), p(self._var),
                                                           p(self._std), _lib.current_stream()))

    def inverse(self, y):
        return y * (self._std + self.eps) + self._mean

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)
        if prefix + "count" in state_dict:
            self._count = int(state_dict[prefix + "count"])


class ActorCritic(nn.Module):
    is_recurrent = False

    def __init__(self, obs, obs_groups, num_actions, actor_obs_normalization=False, critic_obs_normalization=False,
                 actor_hidden_dims=[256, 256, 256], critic_hidden_dims=[256, 256, 256], activation="elu",
                 init_noise_std=1.0, noise_std_type: str = "scalar", device=None, precision: int = 2, **kwargs):
        if kwargs:
            print("ActorCritic.__init__ got unexpected arguments, which will be ignored: " + str(list(kwargs.keys())))
        super().__init__()
        if activation.lower() != "elu":
            raise ValueError("isaaclab_b200 ActorCritic implements the ELU MLP used by the supported configs")
        if noise_std_type != "scalar":
            raise ValueError("isaaclab_b200 ActorCritic supports noise_std_type='scalar'")
        self.obs_groups = obs_groups
        num_actor_obs = sum(obs[g].shape[-1] for g in obs_groups["policy"])
        num_critic_obs = sum(obs[g].shape[-1] for g in obs_groups["critic"])
        for g in list(obs_groups["policy"]) + list(obs_groups["critic"]):
            assert len(obs[g].shape) == 2, "The ActorCritic module only supports 1D observations."
        if device is None:
            device = next(iter(obs.values())).device if hasattr(obs, "values") else "cuda"
        self.actor = _mlp(num_actor_obs, num_actions, actor_hidden_dims)
        self.critic = _mlp(num_critic_obs, 1, critic_hidden_dims)
        self.actor_obs_normalization = bool(actor_obs_normalization)
        self.critic_obs_normalization = bool(critic_obs_normalization)
        self.actor_obs_normalizer = (EmpiricalNormalization(num_actor_obs, device=device)
                                     if actor_obs_normalization else nn.Identity())
        self.critic_obs_normalizer = (EmpiricalNormalization(num_critic_obs, device=device)
                                      if critic_obs_normalization else nn.Identity())
        self.noise_std_type = noise_std_type
        self.std = nn.Parameter(init_noise_std * torch.ones(num_actions))
        self.num_actions = num_actions
        self.precision = precision
        self._flatten(torch.device(device))
        self._a = _Net(self.actor)
        self._c = _Net(self.critic)
        self._mean = None
        self._value = None
        self._sigma_cache = {}

    # -- flat parameter / gradient storage ---------------------------------------------------------
    def _flatten(self, device):
        """All parameters as views of ONE flat buffer.  The first layer of each MLP reads the observation row (235 wide for
        the rough tasks): when that width is not a multiple of 4 its weight is stored in a block padded to the next
        multiple -- the parameter stays a [out, in] (strided) view with the reference's shape and state_dict entry, the pad
        columns are zero and receive zero gradients, and the layer's GEMMs see 16-byte aligned rows (TMA / vector copies
        instead of the 4-byte path: profiles/r2_notes.md)."""
        params = list(self.parameters())
        padded = {id(m[0].weight) for m in (getattr(self, "actor", None), getattr(self, "critic", None))
                  if isinstance(m, nn.Sequential) and m[0].in_features % 4 != 0}
        blocks = []
        for p in params:
            kp = (p.shape[1] + 3) // 4 * 4 if id(p) in padded else None
            blocks.append((p.shape[0] * kp) if kp else p.numel())
        n = sum(blocks)
        self.flat_param = torch.zeros(n, device=device)
        # gradient buffer carries 8 extra floats (loss statistics incl. the KL) so one all-reduce moves both
        self.flat_grad_ext = torch.zeros(n + 8, device=device)
        self.flat_grad = self.flat_grad_ext[:n]
        self.stats = self.flat_grad_ext[n:]
        off = 0
        self._views = []
        for p, k in zip(params, blocks):
            if id(p) in padded:
                kp = k // p.shape[0]
                view = self.flat_param[off:off + k].view(p.shape[0], kp)[:, :p.shape[1]]
                view.copy_(p.data)
                p.data = view
                p.grad = self.flat_grad[off:off + k].view(p.shape[0], kp)[:, :p.shape[1]]
            else:
                self.flat_param[off:off + k].copy_(p.data.reshape(-1))
                p.data = self.flat_param[off:off + k].view(p.shape)
                p.grad = self.flat_grad[off:off + k].view(p.shape)
            p.requires_grad_(False)
            self._views.append((off, k))
            off += k
        self.num_params = n
        self._std_off = self._views[[id(p) for p in params].index(id(self.std))][0]

    def _apply(self, fn, recurse=True):  # keep the flat views alive across .to()/.cuda() calls
        probe = fn(self.flat_param)
        if probe.device != self.flat_param.device or probe.dtype != self.flat_param.dtype:
            raise RuntimeError("isaaclab_b200.ActorCritic lives on one CUDA device in fp32; construct it there")
        return self

    def load_state_dict(self, state_dict, strict=True):
        own = dict(self.named_parameters())
        own.update(dict(self.named_buffers()))
        missing = [k for k in own if k not in state_dict]
        unexpected = [k for k in state_dict if k not in own]
        if strict and (missing or unexpected):
            raise RuntimeError(f"state_dict mismatch: missing {missing}, unexpected {unexpected}")
        for k, p in own.items():
            if k in state_dict:
                p.data.copy_(state_dict[k].to(p.device))
        for m in (self.actor_obs_normalizer, self.critic_obs_normalizer):
            if isinstance(m, EmpiricalNormalization):
                m._count = int(m.count.item())
        return True  # training resumes (RSL/modules/actor_critic.py:166-168)

    # -- rsl_rl surface ----------------------------------------------------------------------------
    def reset(self, dones=None):
        pass

    def forward(self):
        raise NotImplementedError

    def get_actor_obs(self, obs):
        ts = [obs[g] for g in self.obs_groups["policy"]]
        return ts[0] if len(ts) == 1 else torch.cat(ts, dim=-1)

    def get_critic_obs(self, obs):
        ts = [obs[g] for g in self.obs_groups["critic"]]
        return ts[0] if len(ts) == 1 else torch.cat(ts, dim=-1)

    def actor_input(self, obs):
        """Actor observation set, normalised when enabled (actor_critic.py:115-117)."""
        x = self.get_actor_obs(obs)
        if self.actor_obs_normalization:
            return self.actor_obs_normalizer(x.contiguous().float())
        return x if getattr(x, "_b200_row_padded", False) else x.contiguous().float()

    def critic_input(self, obs):
        x = self.get_critic_obs(obs)
        if self.critic_obs_normalization:
            return self.critic_obs_normalizer(x.contiguous().float())
        return x if getattr(x, "_b200_row_padded", False) else x.contiguous().float()

    @property
    def action_mean(self):
        return self._mean

    @property
    def action_std(self):
        M = self._mean.shape[0]
        return self.std.detach().expand(M, -1)

    @property
    def entropy(self):
        e = (0.5 + 0.5 * math.log(2 * math.pi) + torch.log(self.std.detach())).sum()
        return e.expand(self._mean.shape[0])

    def update_distribution(self, obs):
        self._mean = self._a.forward(obs.contiguous().float(), self.precision)

    def act(self, obs, eps=None, **kwargs):
        """Normal(actor(obs), std).sample().  `eps` (standard-normal draws, [M,A]) may be passed explicitly
        (parity mode); otherwise drawn from torch's CUDA generator."""
        x = self.actor_input(obs)
        self.update_distribution(x)
        M, A = self._mean.shape
        if eps is None:
            eps = torch.randn(M, A, device=x.device)
        actions = torch.empty(M, A, device=x.device)
        self._log_prob = torch.empty(M, device=x.device)
        _lib.check(_lib.lib().b200lab_sample_actions(_lib.ptr(self._mean), _lib.ptr(self.std), _lib.ptr(eps),
                                                    _lib.ptr(actions), _lib.ptr(self._log_prob), None, M, A,
                                                    _lib.current_stream()))
        self._last_actions = actions
        return actions

    def act_inference(self, obs):
        # public entry points return fresh tensors, as the reference does; the per-batch-size activation buffers
        # (`_Net.buffers`) are reused by the next forward of the same batch size
        return self._a.forward(self.actor_input(obs), self.precision).clone()

    def evaluate(self, obs, **kwargs):
        return self._evaluate(obs).clone()

    def _evaluate(self, obs):
        """Zero-copy critic forward for PPO.act / compute_returns: the result is consumed (stored in the rollout slab)
        before the next critic forward of the same batch size."""
        self._value = self._c.forward(self.critic_input(obs), self.precision)
        return self._value

    def get_actions_log_prob(self, actions):
        if actions is getattr(self, "_last_actions", None):
            return self._log_prob
        mu, sigma = self._mean, self.std.detach()
        return (-((actions - mu) ** 2) / (2 * sigma ** 2) - torch.log(sigma) - 0.5 * math.log(2 * math.pi)).sum(-1)

    def update_normalization(self, obs):
        if self.actor_obs_normalization:
            self.actor_obs_normalizer.update(self.get_actor_obs(obs))
        if self.critic_obs_normalization:
            self.critic_obs_normalizer.update(self.get_critic_obs(obs))


def unpad_trajectories(trajectories, masks):
    """Inverse of `split_and_pad_trajectories` (RSL/utils/utils.py:133-141): [T, n_traj, H] + masks [T, n_traj] ->
    [T, n_envs, H], the valid steps in env-major order."""
    T = trajectories.shape[0]
    tr, t = masks.t().nonzero(as_tuple=True)           # valid (trajectory, step) pairs, trajectory-major = env-major
    return trajectories[t, tr].view(-1, T, trajectories.shape[-1]).transpose(1, 0)


class Memory(nn.Module):
    """RSL/networks/memory.py:14-70 -- the recurrent core of `ActorCriticRecurrent` (GRU or LSTM, `num_layers` deep).

    `self.rnn` is a torch.nn.LSTM / GRU used as the PARAMETER CONTAINER only (same state_dict keys as the reference:
    `rnn.weight_ih_l0`, ...); the arithmetic is explicit: per layer ONE GEMM projects every time step's input, then each
    step is one h.W_hh^T GEMM + one cell kernel (b200lab_lstm_cell_fwd / b200lab_gru_cell_fwd).  Batch mode keeps what
    back-propagation through time needs; `backward` walks the steps in reverse with the cell-derivative kernels and leaves
    dW_ih / dW_hh / db in the parameters' `.grad` views (two GEMMs per layer over all steps at once)."""

    def __init__(self, input_size, type="lstm", num_layers=1, hidden_size=256, device="cuda", precision=2):
        super().__init__()
        self.is_lstm = type.lower() != "gru"
        rnn_cls = nn.LSTM if self.is_lstm else nn.GRU
        self.rnn = rnn_cls(input_size=input_size, hidden_size=hidden_size, num_layers=num_layers).to(device)
        self.input_size, self.hidden_size, self.num_layers = int(input_size), int(hidden_size), int(num_layers)
        self.G = 4 if self.is_lstm else 3
        self.hidden_states = None
        self.precision = precision
        self._saved = None
        self._zero_bias = None

    def _w(self, l):
        r = self.rnn
        return (getattr(r, f"weight_ih_l{l}"), getattr(r, f"weight_hh_l{l}"), getattr(r, f"bias_ih_l{l}"),
                getattr(r, f"bias_hh_l{l}"))

    def _linear(self, x, w, b, out):
        M, K = x.shape
        _lib.check(_lib.lib().b200lab_linear_fwd(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(out), M, K, w.shape[0], 0,
                                                 self.precision, _lib.current_stream()))
        return out

    def _run(self, x, h0, c0, save):
        """x [T, B, I]; h0 (and c0) [L, B, H].  Returns out [T, B, H] and the final states [L, B, H]."""
        L, s, p = _lib.lib(), _lib.current_stream(), _lib.ptr
        T, B, _ = x.shape
        H, G, dev = self.hidden_size, self.G, x.device
        hn = torch.empty(self.num_layers, B, H, device=dev)
        cn = torch.empty(self.num_layers, B, H, device=dev) if self.is_lstm else None
        saved = []
        inp = x.contiguous()
        gh = torch.empty(B, G * H, device=dev)
        for l in range(self.num_layers):
            w_ih, w_hh, b_ih, b_hh = self._w(l)
            gi = self._linear(inp.view(T * B, -1), w_ih, b_ih, torch.empty(T * B, G * H, device=dev)).view(T, B, G * H)
            hall = torch.empty(T + 1, B, H, device=dev)
            hall[0].copy_(h0[l])
            call = None
            if self.is_lstm:
                call = torch.empty(T + 1, B, H, device=dev)
                call[0].copy_(c0[l])
            act = torch.empty(T, B, 4 * H, device=dev) if save else None
            for t in range(T):
                self._linear(hall[t], w_hh, b_hh, gh)
                a_t = p(act[t]) if save else None
                if self.is_lstm:
                    _lib.check(L.b200lab_lstm_cell_fwd(p(gi[t]), p(gh), p(call[t]), p(hall[t + 1]), p(call[t + 1]), a_t, B, H, s))
                else:
                    _lib.check(L.b200lab_gru_cell_fwd(p(gi[t]), p(gh), p(hall[t]), p(hall[t + 1]), a_t, B, H, s))
            hn[l].copy_(hall[T])
            if self.is_lstm:
                cn[l].copy_(call[T])
            if save:
                saved.append(dict(inp=inp, hall=hall, call=call, act=act))
            inp = hall[1:]
        self._saved = saved if save else None
        return inp, hn, cn

    def forward(self, input, masks=None, hidden_states=None):
        batch_mode = masks is not None
        if batch_mode:   # policy update: padded trajectories [T, n_traj, I] from the saved hidden states (memory.py:29-34)
            if hidden_states is None:
                raise ValueError("Hidden states not passed to memory module during policy update")
            h0, c0 = (hidden_states if self.is_lstm else (hidden_states, None))
            out, _, _ = self._run(input, h0.contiguous(), c0.contiguous() if c0 is not None else None, save=True)
            self._masks = masks
            self._padded_shape = out.shape
            return unpad_trajectories(out, masks)
        # inference / rollout: one step from the hidden states of the last step (memory.py:36-37)
        x = input.unsqueeze(0)
        B = x.shape[1]
        if self.hidden_states is None:
            z = torch.zeros(self.num_layers, B, self.hidden_size, device=x.device)
            h0, c0 = z, (z.clone() if self.is_lstm else None)
        else:
            h0, c0 = (self.hidden_states if self.is_lstm else (self.hidden_states, None))
        out, hn, cn = self._run(x, h0, c0, save=False)
        self.hidden_states = (hn, cn) if self.is_lstm else hn
        return out

    def backward(self, d_out):
        """d_out: gradient of the UNPADDED output of the last batch-mode forward, [T, n_envs, H].  Padded steps receive
        zero gradient (their outputs were dropped by `unpad_trajectories`), which is what autograd gives the reference."""
        L, s, p = _lib.lib(), _lib.current_stream(), _lib.ptr
        masks = self._masks
        T, B, H = self._padded_shape
        G, dev = self.G, d_out.device
        tr, t_idx = masks.t().nonzero(as_tuple=True)
        d_pad = torch.zeros(T, B, H, device=dev)
        d_pad[t_idx, tr] = d_out.transpose(1, 0).reshape(-1, H)
        if self._zero_bias is None:
            self._zero_bias = torch.zeros(H, device=dev)
        dh_top = d_pad
        for l in reversed(range(self.num_layers)):
            sv = self._saved[l]
            w_ih, w_hh, b_ih, b_hh = self._w(l)
            w_hh_t = w_hh.t().contiguous()                       # [H, G*H]: dh_{t-1} += dgates_t . W_hh as a linear_fwd
            dgi = torch.empty(T, B, G * H, device=dev)
            dgh = dgi if self.is_lstm else torch.empty(T, B, G * H, device=dev)
            dh_rec = torch.empty(B, H, device=dev)
            carry = torch.zeros(B, H, device=dev)                # LSTM: dL/dc; GRU: the z * h_{t-1} path of dL/dh
            for t in reversed(range(T)):
                rec = p(dh_rec) if t < T - 1 else None
                if self.is_lstm:
                    _lib.check(L.b200lab_lstm_cell_bwd(p(sv["act"][t]), p(sv["call"][t]), p(sv["call"][t + 1]), p(dh_top[t]), rec,
                                                       p(carry), p(dgi[t]), B, H, s))
                else:
                    _lib.check(L.b200lab_gru_cell_bwd(p(sv["act"][t]), p(sv["hall"][t]), p(dh_top[t]), rec, p(carry), p(carry),
                                                      p(dgi[t]), p(dgh[t]), B, H, s))
                if t > 0:
                    self._linear(dgh[t], w_hh_t, self._zero_bias, dh_rec)
            inp = sv["inp"].view(T * B, -1)
            K = inp.shape[1]
            ws = max(L.b200lab_linear_bwd_workspace_bytes(T * B, K, G * H), L.b200lab_linear_bwd_workspace_bytes(T * B, H, G * H))
            work = torch.empty((ws + 3) // 4, device=dev)
            dx = torch.empty(T, B, K, device=dev) if l > 0 else None
            _lib.check(L.b200lab_linear_bwd(p(inp), p(w_ih), None, p(dgi), None, p(dx), p(w_ih.grad), p(b_ih.grad), T * B, K,
                                            G * H, 0, self.precision, p(work), s))
            hprev = sv["hall"][:T].view(T * B, H)
            _lib.check(L.b200lab_linear_bwd(p(hprev), p(w_hh), None, p(dgh), None, None, p(w_hh.grad), p(b_hh.grad), T * B, H,
                                            G * H, 0, self.precision, p(work), s))
            dh_top = dx
        self._saved = None

    def reset(self, dones=None, hidden_states=None):
        if dones is None:   # reset all hidden states (memory.py:41-45)
            self.hidden_states = hidden_states
        elif self.hidden_states is not None:   # zero the rows of the finished envs, on the device, no sync
            if hidden_states is not None:
                raise NotImplementedError("Resetting hidden states of done environments with custom hidden states is not implemented")
            m = (dones == 1).view(1, -1, 1)
            for hs in (self.hidden_states if isinstance(self.hidden_states, tuple) else (self.hidden_states,)):
                hs.masked_fill_(m, 0.0)

    def detach_hidden_states(self, dones=None):
        """No autograd graph hangs off the hidden states on this path (memory.py:57-70 is a no-op here)."""


class ActorCriticRecurrent(ActorCritic):
    """RSL/modules/actor_critic_recurrent.py:17-190: `Memory` (LSTM / GRU) in front of the actor and critic MLPs.  Same
    constructor kwargs, method surface and state_dict keys (`memory_a.rnn.*`, `actor.*`, `memory_c.rnn.*`, `critic.*`,
    `std`); parameters live in the same flat buffers as `ActorCritic`'s."""
    is_recurrent = True

    def __init__(self, obs, obs_groups, num_actions, actor_obs_normalization=False, critic_obs_normalization=False,
                 actor_hidden_dims=[256, 256, 256], critic_hidden_dims=[256, 256, 256], activation="elu",
                 init_noise_std=1.0, noise_std_type: str = "scalar", rnn_type="lstm", rnn_hidden_dim=256, rnn_num_layers=1,
                 device=None, precision: int = 2, **kwargs):
        if "rnn_hidden_size" in kwargs:   # deprecated spelling (actor_critic_recurrent.py:36-43)
            import warnings
            warnings.warn("The argument `rnn_hidden_size` is deprecated and will be removed in a future version. "
                          "Please use `rnn_hidden_dim` instead.", DeprecationWarning)
            size = kwargs.pop("rnn_hidden_size")
            if rnn_hidden_dim == 256:
                rnn_hidden_dim = size
        if kwargs:
            print("ActorCriticRecurrent.__init__ got unexpected arguments, which will be ignored: " + str(kwargs.keys()))
        nn.Module.__init__(self)
        if activation.lower() != "elu":
            raise ValueError("isaaclab_b200 ActorCriticRecurrent implements the ELU MLP used by the supported configs")
        if noise_std_type != "scalar":
            raise ValueError("isaaclab_b200 ActorCriticRecurrent supports noise_std_type='scalar'")
        self.obs_groups = obs_groups
        for g in list(obs_groups["policy"]) + list(obs_groups["critic"]):
            assert len(obs[g].shape) == 2, "The ActorCriticRecurrent module only supports 1D observations."
        num_actor_obs = sum(obs[g].shape[-1] for g in obs_groups["policy"])
        num_critic_obs = sum(obs[g].shape[-1] for g in obs_groups["critic"])
        if device is None:
            device = next(iter(obs.values())).device if hasattr(obs, "values") else "cuda"
        # registration order = the reference's, so named_parameters() / optimizer state indices line up
        self.memory_a = Memory(num_actor_obs, type=rnn_type, num_layers=rnn_num_layers, hidden_size=rnn_hidden_dim,
                               device=device, precision=precision)
        self.actor = _mlp(rnn_hidden_dim, num_actions, actor_hidden_dims)
        self.actor_obs_normalization = bool(actor_obs_normalization)
        self.actor_obs_normalizer = (EmpiricalNormalization(num_actor_obs, device=device)
                                     if actor_obs_normalization else nn.Identity())
        self.memory_c = Memory(num_critic_obs, type=rnn_type, num_layers=rnn_num_layers, hidden_size=rnn_hidden_dim,
                               device=device, precision=precision)
        self.critic = _mlp(rnn_hidden_dim, 1, critic_hidden_dims)
        self.critic_obs_normalization = bool(critic_obs_normalization)
        self.critic_obs_normalizer = (EmpiricalNormalization(num_critic_obs, device=device)
                                      if critic_obs_normalization else nn.Identity())
        self.noise_std_type = noise_std_type
        self.std = nn.Parameter(init_noise_std * torch.ones(num_actions))
        self.num_actions = num_actions
        self.precision = precision
        self._flatten(torch.device(device))
        self._a = _Net(self.actor)
        self._c = _Net(self.critic)
        self._mean = None
        self._value = None
        self._sigma_cache = {}

    def reset(self, dones=None):
        self.memory_a.reset(dones)
        self.memory_c.reset(dones)

    def get_hidden_states(self):
        return self.memory_a.hidden_states, self.memory_c.hidden_states

    def _normalized(self, x, normalizer, on):
        x = x.contiguous().float()
        if not on:
            return x
        return normalizer(x.view(-1, x.shape[-1])).view(x.shape)   # padded trajectories are [T, n_traj, D]

    def actor_features(self, obs, masks=None, hidden_states=None):
        """Normalised actor observations through `memory_a`: [B, H] in rollout mode, [T * n_envs, H] in batch mode."""
        x = self._normalized(self.get_actor_obs(obs), self.actor_obs_normalizer, self.actor_obs_normalization)
        out = self.memory_a(x, masks, hidden_states)
        return out.squeeze(0) if masks is None else out.reshape(-1, out.shape[-1])

    def critic_features(self, obs, masks=None, hidden_states=None):
        x = self._normalized(self.get_critic_obs(obs), self.critic_obs_normalizer, self.critic_obs_normalization)
        out = self.memory_c(x, masks, hidden_states)
        return out.squeeze(0) if masks is None else out.reshape(-1, out.shape[-1])

    def act(self, obs, masks=None, hidden_states=None, eps=None, **kwargs):
        f = self.actor_features(obs, masks, hidden_states).contiguous()
        self._mean = self._a.forward(f, self.precision)
        M, A = self._mean.shape
        if eps is None:
            eps = torch.randn(M, A, device=f.device)
        actions = torch.empty(M, A, device=f.device)
        self._log_prob = torch.empty(M, device=f.device)
        _lib.check(_lib.lib().b200lab_sample_actions(_lib.ptr(self._mean), _lib.ptr(self.std), _lib.ptr(eps),
                                                    _lib.ptr(actions), _lib.ptr(self._log_prob), None, M, A,
                                                    _lib.current_stream()))
        self._last_actions = actions
        return actions

    def act_inference(self, obs):
        return self._a.forward(self.actor_features(obs).contiguous(), self.precision).clone()

    def evaluate(self, obs, masks=None, hidden_states=None, **kwargs):
        return self._evaluate(obs, masks, hidden_states).clone()

    def _evaluate(self, obs, masks=None, hidden_states=None):
        self._value = self._c.forward(self.critic_features(obs, masks, hidden_states).contiguous(), self.precision)
        return self._value
