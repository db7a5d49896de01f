"""The quaternion helpers user terms call (`isaaclab.utils.math`, LAB/utils/math.py) as plain torch on the env's device.
Off the fused path: the kernels restate these formulas themselves (csrc/common.cuh); these exist so that a Python term
written against the reference's `math_utils` runs unchanged.  Quaternions are (w, x, y, z)."""
from __future__ import annotations

import math

import torch


def normalize(x: torch.Tensor, eps: float = 1e-9) -> torch.Tensor:
    return x / x.norm(p=2, dim=-1).clamp(min=eps, max=None).unsqueeze(-1)   # math.py:82


def wrap_to_pi(angles: torch.Tensor) -> torch.Tensor:
    wrapped = (angles + math.pi) % (2 * math.pi)   # math.py:96-117
    return torch.where((wrapped == 0) & (angles > 0), math.pi, wrapped - math.pi)


def quat_apply(quat: torch.Tensor, vec: torch.Tensor) -> torch.Tensor:
    shape = vec.shape   # math.py:546-564
    quat, vec = quat.reshape(-1, 4), vec.reshape(-1, 3)
    xyz = quat[:, 1:]
    t = xyz.cross(vec, dim=-1) * 2
    return (vec + quat[:, 0:1] * t + xyz.cross(t, dim=-1)).view(shape)


def quat_rotate_inverse(q: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    q_w, q_vec = q[..., 0], q[..., 1:]   # math.py:606-625
    a = v * (2.0 * q_w**2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * (q_vec * v).sum(-1, keepdim=True) * 2.0
    return a - b + c


def quat_rotate(q: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    q_w, q_vec = q[..., 0], q[..., 1:]
    a = v * (2.0 * q_w**2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * (q_vec * v).sum(-1, keepdim=True) * 2.0
    return a + b + c


def yaw_quat(quat: torch.Tensor) -> torch.Tensor:
    shape = quat.shape   # math.py:522-542
    q = quat.reshape(-1, 4)
    qw, qx, qy, qz = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    yaw = torch.atan2(2 * (qw * qz + qx * qy), 1 - 2 * (qy * qy + qz * qz))
    out = torch.zeros_like(q)
    out[:, 3] = torch.sin(yaw / 2)
    out[:, 0] = torch.cos(yaw / 2)
    return normalize(out).view(shape)


def quat_apply_yaw(quat: torch.Tensor, vec: torch.Tensor) -> torch.Tensor:
    return quat_apply(yaw_quat(quat), vec)   # math.py:568-579
