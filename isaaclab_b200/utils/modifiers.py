"""Observation modifiers with the reference's names (LAB/utils/modifiers/modifier.py:23-66, modifier_cfg.py:17-36):
`ModifierCfg(func=scale|clip|bias, params={...})` entries of `ObservationTermCfg.modifiers`.

On the fused path the three function modifiers are compiled into the column program of the term (applied before the
noise, observation_manager.py:310-312); called directly they are plain torch functions, which is how user terms evaluated
in Python get them."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable

import torch


def scale(data: torch.Tensor, multiplier: float) -> torch.Tensor:
    return data * multiplier


def clip(data: torch.Tensor, bounds: tuple) -> torch.Tensor:
    """bounds = (min | None, max | None)"""
    return data.clip(min=bounds[0], max=bounds[1])


def bias(data: torch.Tensor, value: float) -> torch.Tensor:
    return data + value


scale._b200_modifier, clip._b200_modifier, bias._b200_modifier = "scale", "clip", "bias"


@dataclass
class ModifierCfg:
    func: Callable[..., torch.Tensor] = None
    params: dict[str, Any] = field(default_factory=dict)


# ------------------------------------------------------------------------------------------------------------------
# Class modifiers (LAB/utils/modifiers/modifier_base.py, modifier.py:69-260, modifier_cfg.py:37-78): stateful, one instance per
# observation term, constructed with the term's output shape, `reset(env_ids)` when envs reset.  They keep their state in
# torch tensors; a term that carries one is evaluated through the EXTERNAL column path (`envs/compile.py`), whatever its
# function -- fused identities included (they are evaluated with a one-term program, then modified here).
# ------------------------------------------------------------------------------------------------------------------
class ModifierBase:
    def __init__(self, cfg, data_dim: tuple, device: str):
        self._cfg, self._data_dim, self._device = cfg, tuple(data_dim), device

    def reset(self, env_ids=None):
        raise NotImplementedError

    def __call__(self, data: torch.Tensor) -> torch.Tensor:
        raise NotImplementedError


class DigitalFilter(ModifierBase):
    """y_i = sum_j B[j] x_{i-j} - sum_{j>=1} A[j-1] y_{i-j}  (linear difference equation; `A` lists the feedback coefficients
    of the PREVIOUS outputs, `B` the feed-forward ones starting at the current input).  Histories are rings addressed by a
    write index -- no roll."""

    def __init__(self, cfg, data_dim, device):
        super().__init__(cfg, data_dim, device)
        if cfg.A is None or cfg.B is None:
            raise ValueError("Digital filter coefficients A and B must not be None. Please provide valid coefficients.")
        self.A = torch.tensor(list(cfg.A), dtype=torch.float32, device=device)
        self.B = torch.tensor(list(cfg.B), dtype=torch.float32, device=device)
        self.x_hist = torch.zeros(self._data_dim + (self.B.numel(),), device=device)
        self.y_hist = torch.zeros(self._data_dim + (max(self.A.numel(), 1),), device=device)
        self._px = self._py = 0

    def reset(self, env_ids=None):
        ids = slice(None) if env_ids is None else env_ids
        self.x_hist[ids] = 0.0
        self.y_hist[ids] = 0.0

    def __call__(self, data):
        nb, na = self.B.numel(), self.A.numel()
        self._px = (self._px + 1) % nb
        self.x_hist[..., self._px] = data
        order_x = [(self._px - j) % nb for j in range(nb)]              # newest first
        y = (self.x_hist[..., order_x] * self.B).sum(-1)
        if na:
            order_y = [(self._py - j) % na for j in range(na)]
            y = y - (self.y_hist[..., order_y] * self.A).sum(-1)
            self._py = (self._py + 1) % na
            self.y_hist[..., self._py] = y
        return y


class Integrator(ModifierBase):
    """Trapezoidal running integral: integral += (x_i + x_{i-1}) / 2 * dt."""

    def __init__(self, cfg, data_dim, device):
        super().__init__(cfg, data_dim, device)
        self.integral = torch.zeros(self._data_dim, device=device)
        self.y_prev = torch.zeros(self._data_dim, device=device)

    def reset(self, env_ids=None):
        ids = slice(None) if env_ids is None else env_ids
        self.integral[ids] = 0.0
        self.y_prev[ids] = 0.0

    def __call__(self, data):
        self.integral += (data + self.y_prev) / 2 * self._cfg.dt
        self.y_prev[:] = data
        return self.integral


@dataclass
class DigitalFilterCfg(ModifierCfg):
    func: Any = DigitalFilter
    A: list = None
    B: list = None


@dataclass
class IntegratorCfg(ModifierCfg):
    func: Any = Integrator
    dt: float = None
