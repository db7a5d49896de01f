"""Oracle restatement of the math helpers the hot path uses (test infrastructure only).

LAB = /root/reference/source/isaaclab/isaaclab ; every function cites LAB/utils/math.py.
"""
from __future__ import annotations

import torch


def wrap_to_pi(angles):
    """math.py:96-117: wrap to [-pi, pi], keeping +pi for positive odd multiples of pi."""
    wrapped = (angles + torch.pi) % (2 * torch.pi)
    return torch.where((wrapped == 0) & (angles > 0), torch.pi, wrapped - torch.pi)


def normalize(x, eps=1e-9):
    """math.py:82-93."""
    return x / x.norm(p=2, dim=-1).clamp(min=eps, max=None).unsqueeze(-1)


def quat_rotate_inverse(q, v):
    """math.py:606-625 (q = wxyz): a - b + c."""
    q_w = q[..., 0]
    q_vec = q[..., 1:]
    a = v * (2.0 * q_w**2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * (q_vec * v).sum(-1, keepdim=True) * 2.0
    return a - b + c


def quat_apply(quat, vec):
    """math.py:546-564: vec + w*t + xyz x t with t = 2 * (xyz x vec)."""
    xyz = quat[..., 1:]
    t = torch.cross(xyz, vec, dim=-1) * 2
    return vec + quat[..., 0:1] * t + torch.cross(xyz, t, dim=-1)


def yaw_quat(quat):
    """math.py:522-542."""
    qw, qx, qy, qz = quat[..., 0], quat[..., 1], quat[..., 2], quat[..., 3]
    yaw = torch.atan2(2 * (qw * qz + qx * qy), 1 - 2 * (qy * qy + qz * qz))
    out = torch.zeros_like(quat)
    out[..., 3] = torch.sin(yaw / 2)
    out[..., 0] = torch.cos(yaw / 2)
    return normalize(out)


def quat_apply_yaw(quat, vec):
    """math.py:568-579."""
    return quat_apply(yaw_quat(quat), vec)


def quat_from_euler_xyz(roll, pitch, yaw):
    """math.py:252-280."""
    cy, sy = torch.cos(yaw * 0.5), torch.sin(yaw * 0.5)
    cr, sr = torch.cos(roll * 0.5), torch.sin(roll * 0.5)
    cp, sp = torch.cos(pitch * 0.5), torch.sin(pitch * 0.5)
    qw = cy * cr * cp + sy * sr * sp
    qx = cy * sr * cp - sy * cr * sp
    qy = cy * cr * sp + sy * sr * cp
    qz = sy * cr * cp - cy * sr * sp
    return torch.stack([qw, qx, qy, qz], dim=-1)


def quat_mul(q1, q2):
    """math.py:464-500."""
    w1, x1, y1, z1 = q1[..., 0], q1[..., 1], q1[..., 2], q1[..., 3]
    w2, x2, y2, z2 = q2[..., 0], q2[..., 1], q2[..., 2], q2[..., 3]
    ww = (z1 + x1) * (x2 + y2)
    yy = (w1 - y1) * (w2 + z2)
    zz = (w1 + y1) * (w2 - z2)
    xx = ww + yy + zz
    qq = 0.5 * (xx + (z1 - x1) * (x2 - y2))
    w = qq - ww + (z1 - y1) * (y2 - z2)
    x = qq - xx + (x1 + w1) * (x2 + w2)
    y = qq - yy + (w1 - x1) * (y2 + z2)
    z = qq - zz + (z1 + y1) * (w2 - x2)
    return torch.stack([w, x, y, z], dim=-1)


def scale_uniform(u, lower, upper):
    """math.py:1313-1331 sample_uniform with the [0,1) draw `u` made explicit: u * (upper - lower) + lower."""
    return u * (upper - lower) + lower
