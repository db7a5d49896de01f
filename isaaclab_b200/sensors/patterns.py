"""Ray patterns of the ray-caster sensor with the reference's names, cfg fields and ray order
(LAB/sensors/ray_caster/patterns/patterns.py:16-180, patterns_cfg.py:21-220): `grid_pattern` (the height scanner),
`lidar_pattern`, `bpearl_pattern`, `pinhole_camera_pattern`.  Each returns (ray_starts, ray_directions) in the sensor frame;
they are evaluated once, on the host, when the env is built -- the rays then live in the column program / the ray tables of
the kernels (csrc/mdp.cu, csrc/raycast.cu)."""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Sequence

import torch

from .ray_caster import grid_pattern as _grid_numpy


def grid_pattern(cfg, device="cpu"):
    s, d = _grid_numpy(cfg.size, cfg.resolution, cfg.direction, cfg.ordering)
    return torch.as_tensor(s, device=device), torch.as_tensor(d, device=device)


def lidar_pattern(cfg, device="cpu"):
    """channels x horizontal samples, channel-major; a full 360 degree sweep drops its duplicate last column
    (patterns.py:136-180)."""
    v = torch.deg2rad(torch.linspace(cfg.vertical_fov_range[0], cfg.vertical_fov_range[1], cfg.channels))
    lo, hi = cfg.horizontal_fov_range
    n_h = math.ceil((hi - lo) / cfg.horizontal_res)
    h = torch.linspace(lo, hi, n_h)
    if abs(abs(lo - hi) - 360.0) < 1e-6:
        h = h[:-1]
    h = torch.deg2rad(h)
    va, ha = v.unsqueeze(1), h.unsqueeze(0)                      # [C, 1], [1, Hs]
    dirs = torch.stack([(torch.cos(va) * torch.cos(ha)), (torch.cos(va) * torch.sin(ha)),
                        torch.sin(va).expand(-1, h.numel())], dim=-1).reshape(-1, 3).to(device)
    return torch.zeros_like(dirs), dirs


def bpearl_pattern(cfg, device="cpu"):
    """RS-Bpearl: yaw-major over the listed vertical ray angles, directions pointing DOWN-and-out (patterns.py:106-133)."""
    h = torch.arange(-cfg.horizontal_fov / 2, cfg.horizontal_fov / 2, cfg.horizontal_res)
    v = torch.tensor(list(cfg.vertical_ray_angles))
    yaw = torch.deg2rad(h.unsqueeze(1).expand(-1, v.numel()).reshape(-1))       # meshgrid(v, h, "xy") flattened: h outer
    pitch = torch.deg2rad(v.unsqueeze(0).expand(h.numel(), -1).reshape(-1)) + torch.pi / 2
    dirs = -torch.stack([torch.sin(pitch) * torch.cos(yaw), torch.sin(pitch) * torch.sin(yaw), torch.cos(pitch)], dim=1)
    dirs = dirs.to(device)
    return torch.zeros_like(dirs), dirs


def pinhole_camera_pattern(cfg, intrinsic_matrices: torch.Tensor, device="cpu"):
    """One ray through the centre of every pixel, row-major, in the robotics camera frame (x forward, y left, z up);
    `intrinsic_matrices` [N, 3, 3] -> ([N, H*W, 3], [N, H*W, 3])  (patterns.py:61-103)."""
    u = torch.arange(cfg.width, dtype=torch.float32) + 0.5
    v = torch.arange(cfg.height, dtype=torch.float32) + 0.5
    px = torch.stack([u.repeat(cfg.height), v.repeat_interleave(cfg.width), torch.ones(cfg.width * cfg.height)], dim=0)
    cam = torch.matmul(torch.inverse(intrinsic_matrices.float().cpu()), px)          # [N, 3, H*W]: x right, y down, z forward
    rob = torch.stack([cam[:, 2], -cam[:, 0], -cam[:, 1]], dim=1)
    dirs = (rob / torch.norm(rob, dim=1, keepdim=True)).permute(0, 2, 1).contiguous().to(device)
    return torch.zeros_like(dirs), dirs


@dataclass
class PatternBaseCfg:
    func: object = None


@dataclass
class GridPatternCfg(PatternBaseCfg):
    resolution: float = None
    size: tuple = None
    direction: tuple = (0.0, 0.0, -1.0)
    ordering: str = "xy"
    func: object = grid_pattern


@dataclass
class LidarPatternCfg(PatternBaseCfg):
    channels: int = None
    vertical_fov_range: tuple = None
    horizontal_fov_range: tuple = None
    horizontal_res: float = None
    func: object = lidar_pattern


_BPEARL_ANGLES = [89.5, 86.6875, 83.875, 81.0625, 78.25, 75.4375, 72.625, 69.8125, 67.0, 64.1875, 61.375, 58.5625, 55.75,
                  52.9375, 50.125, 47.3125, 44.5, 41.6875, 38.875, 36.0625, 33.25, 30.4375, 27.625, 24.8125, 22, 19.1875,
                  16.375, 13.5625, 10.75, 7.9375, 5.125, 2.3125]


@dataclass
class BpearlPatternCfg(PatternBaseCfg):
    horizontal_fov: float = 360.0
    horizontal_res: float = 10.0
    vertical_ray_angles: Sequence[float] = field(default_factory=lambda: list(_BPEARL_ANGLES))
    func: object = bpearl_pattern


@dataclass
class PinholeCameraPatternCfg(PatternBaseCfg):
    focal_length: float = 24.0
    horizontal_aperture: float = 20.955
    vertical_aperture: float | None = None
    horizontal_aperture_offset: float = 0.0
    vertical_aperture_offset: float = 0.0
    width: int = None
    height: int = None
    func: object = pinhole_camera_pattern

    @classmethod
    def from_intrinsic_matrix(cls, intrinsic_matrix, width, height, focal_length: float = 24.0):
        """patterns_cfg.py:121-173: apertures (and their offsets) from f_x, f_y, c_x, c_y of a row-major 3x3 matrix."""
        f_x, c_x, f_y, c_y = intrinsic_matrix[0], intrinsic_matrix[2], intrinsic_matrix[4], intrinsic_matrix[5]
        h_ap, v_ap = width * focal_length / f_x, height * focal_length / f_y
        return cls(focal_length=focal_length, horizontal_aperture=h_ap, vertical_aperture=v_ap,
                   horizontal_aperture_offset=(c_x - width / 2) / f_x, vertical_aperture_offset=(c_y - height / 2) / f_y,
                   width=width, height=height)
