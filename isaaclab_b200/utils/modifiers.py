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
