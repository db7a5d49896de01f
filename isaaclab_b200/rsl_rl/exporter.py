"""Policy export with the interface of the reference's exporter (RL/rsl_rl/exporter.py:13-45):
`export_policy_as_jit(policy, normalizer, path, filename)` and `export_policy_as_onnx(policy, path, normalizer, filename)`.

The exported artefact is a plain CPU torch module -- actor MLP, optional observation normaliser, optional recurrent memory with
its state kept inside the module and a `reset()` method -- rebuilt from the policy's parameter values, so it carries no
dependency on libb200lab and loads wherever the reference's exported policies load (the deployment side of the boundary)."""
from __future__ import annotations

import os

import torch
from torch import nn


def _plain_mlp(seq: nn.Sequential) -> nn.Sequential:
    """A fresh CPU copy of an MLP whose parameters may be (strided) views of the flat training buffer."""
    layers = []
    for m in seq:
        if isinstance(m, nn.Linear):
            lin = nn.Linear(m.in_features, m.out_features)
            with torch.no_grad():
                lin.weight.copy_(m.weight.detach().cpu())
                lin.bias.copy_(m.bias.detach().cpu())
            layers.append(lin)
        else:
            layers.append(type(m)())
    return nn.Sequential(*layers)


class _Normalizer(nn.Module):
    """(x - mean) / (std + eps) with the statistics frozen (RSL/networks/normalization.py:45-48 in eval mode)."""

    def __init__(self, normalizer):
        super().__init__()
        self.register_buffer("mean", normalizer._mean.detach().cpu().clone())
        self.register_buffer("std", normalizer._std.detach().cpu().clone())
        self.eps = float(normalizer.eps)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return (x - self.mean) / (self.std + self.eps)


class _Feedforward(nn.Module):
    def __init__(self, actor, normalizer):
        super().__init__()
        self.actor, self.normalizer = actor, normalizer

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.actor(self.normalizer(x))

    @torch.jit.export
    def reset(self):
        pass


class _RecurrentLSTM(nn.Module):
    """One env, state inside the module (exporter.py:62-79)."""

    def __init__(self, actor, normalizer, rnn: nn.LSTM):
        super().__init__()
        self.actor, self.normalizer, self.rnn = actor, normalizer, rnn
        self.register_buffer("hidden_state", torch.zeros(rnn.num_layers, 1, rnn.hidden_size))
        self.register_buffer("cell_state", torch.zeros(rnn.num_layers, 1, rnn.hidden_size))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.normalizer(x)
        y, (h, c) = self.rnn(x.unsqueeze(0), (self.hidden_state, self.cell_state))
        self.hidden_state[:] = h
        self.cell_state[:] = c
        return self.actor(y.squeeze(0))

    @torch.jit.export
    def reset(self):
        self.hidden_state[:] = 0.0
        self.cell_state[:] = 0.0


class _RecurrentGRU(nn.Module):
    def __init__(self, actor, normalizer, rnn: nn.GRU):
        super().__init__()
        self.actor, self.normalizer, self.rnn = actor, normalizer, rnn
        self.register_buffer("hidden_state", torch.zeros(rnn.num_layers, 1, rnn.hidden_size))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.normalizer(x)
        y, h = self.rnn(x.unsqueeze(0), self.hidden_state)
        self.hidden_state[:] = h
        return self.actor(y.squeeze(0))

    @torch.jit.export
    def reset(self):
        self.hidden_state[:] = 0.0


def _deployable(policy, normalizer) -> nn.Module:
    if not hasattr(policy, "actor"):
        raise ValueError("Policy does not have an actor/student module.")   # exporter.py:60
    actor = _plain_mlp(policy.actor)
    norm = _Normalizer(normalizer) if normalizer is not None and hasattr(normalizer, "_mean") else nn.Identity()
    if not getattr(policy, "is_recurrent", False):
        return _Feedforward(actor, norm)
    src = policy.memory_a.rnn
    rnn = type(src)(input_size=src.input_size, hidden_size=src.hidden_size, num_layers=src.num_layers)
    rnn.load_state_dict({k: v.detach().cpu().contiguous().clone() for k, v in src.state_dict().items()})
    return (_RecurrentLSTM if isinstance(rnn, nn.LSTM) else _RecurrentGRU)(actor, norm, rnn)


def export_policy_as_jit(policy, normalizer, path: str, filename: str = "policy.pt"):
    """TorchScript file with `forward(obs)` and `reset()` (exporter.py:13-24)."""
    os.makedirs(path, exist_ok=True)
    module = _deployable(policy, normalizer).eval()
    torch.jit.script(module).save(os.path.join(path, filename))


def export_policy_as_onnx(policy, path: str, normalizer=None, filename: str = "policy.onnx", verbose: bool = False):
    """ONNX graph: `obs -> actions`, or `obs, h_in, c_in -> actions, h_out, c_out` for a recurrent policy
    (exporter.py:27-45,138-175)."""
    os.makedirs(path, exist_ok=True)
    module = _deployable(policy, normalizer).eval()
    out = os.path.join(path, filename)
    if isinstance(module, _Feedforward):
        obs = torch.zeros(1, module.actor[0].in_features)
        torch.onnx.export(module, obs, out, export_params=True, opset_version=11, verbose=verbose, input_names=["obs"],
                          output_names=["actions"], dynamic_axes={})
        return

    class _Stateless(nn.Module):   # the ONNX graph takes and returns the memory explicitly
        def __init__(self, m):
            super().__init__()
            self.m = m

        def forward(self, x, h, c):
            x = self.m.normalizer(x)
            if isinstance(self.m.rnn, nn.LSTM):
                y, (h2, c2) = self.m.rnn(x.unsqueeze(0), (h, c))
            else:
                y, h2 = self.m.rnn(x.unsqueeze(0), h)
                c2 = c
            return self.m.actor(y.squeeze(0)), h2, c2

    rnn = module.rnn
    z = torch.zeros(rnn.num_layers, 1, rnn.hidden_size)
    torch.onnx.export(_Stateless(module), (torch.zeros(1, rnn.input_size), z, z.clone()), out, export_params=True,
                      opset_version=11, verbose=verbose, input_names=["obs", "h_in", "c_in"],
                      output_names=["actions", "h_out", "c_out"], dynamic_axes={})
