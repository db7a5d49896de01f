"""`RayCasterCamera` on the libb200lab mesh query (LAB/sensors/ray_caster/ray_caster_camera.py:257-318): a depth camera made
of one ray per pixel.  The pinhole pattern (`patterns.pinhole_camera_pattern`) gives the rays in the camera's robotics
frame; every update rotates them by the FULL camera orientation, casts them with `raycast_mesh` (max_dist 1e6 -- the
clipping happens afterwards, as in the reference, to match the USD camera), and fills

    distance_to_image_plane [N, H, W, 1]   depth along the optical axis (the x component of depth * direction in the camera frame)
    distance_to_camera      [N, H, W, 1]   ray length
    normals                 [N, H, W, 3]   face normal at the hit

with `depth_clipping_behavior` in {"max", "zero", "none"}.  Poses come from the caller (camera position / orientation in the
world, [N, 3] / [N, 4] wxyz): where the reference reads them from a USD prim view, this class takes the tensors."""
from __future__ import annotations

from dataclasses import dataclass, field

import torch

from ..utils import math as M
from . import patterns
from .ray_caster import WarpFreeMesh, raycast_mesh


@dataclass
class RayCasterCameraCfg:
    pattern_cfg: patterns.PinholeCameraPatternCfg = None
    data_types: list = field(default_factory=lambda: ["distance_to_image_plane"])
    max_distance: float = 1.0e6
    depth_clipping_behavior: str = "none"   # "max" | "zero" | "none"


def _quat_inv(q):
    return torch.cat([q[..., :1], -q[..., 1:]], dim=-1) / (q * q).sum(-1, keepdim=True)


class RayCasterCamera:
    UNSUPPORTED_TYPES = {"rgb", "instance_id_segmentation", "instance_id_segmentation_fast", "instance_segmentation",
                         "instance_segmentation_fast", "semantic_segmentation", "skeleton_data", "motion_vectors",
                         "bounding_box_2d_tight", "bounding_box_2d_tight_fast", "bounding_box_2d_loose",
                         "bounding_box_2d_loose_fast", "bounding_box_3d", "bounding_box_3d_fast"}

    def __init__(self, cfg: RayCasterCameraCfg, mesh: WarpFreeMesh, num_envs: int, intrinsic_matrices: torch.Tensor | None = None,
                 device="cuda"):
        bad = set(cfg.data_types) & self.UNSUPPORTED_TYPES
        if bad:   # ray_caster_camera.py:325-337
            raise ValueError(f"RayCasterCamera class does not support the following sensor types: {bad}.")
        if cfg.depth_clipping_behavior not in ("max", "zero", "none"):
            raise ValueError(f"unknown depth_clipping_behavior '{cfg.depth_clipping_behavior}'")
        self.cfg, self.mesh, self.num_envs, self.device = cfg, mesh, int(num_envs), torch.device(device)
        pc = cfg.pattern_cfg
        self.image_shape = (pc.height, pc.width)
        if intrinsic_matrices is None:   # from the apertures (ray_caster_camera.py:391-408)
            v_ap = pc.vertical_aperture if pc.vertical_aperture is not None else pc.horizontal_aperture * pc.height / pc.width
            f_x, f_y = pc.width * pc.focal_length / pc.horizontal_aperture, pc.height * pc.focal_length / v_ap
            c_x = pc.horizontal_aperture_offset * f_x + pc.width / 2
            c_y = pc.vertical_aperture_offset * f_y + pc.height / 2
            k = torch.tensor([[f_x, 0.0, c_x], [0.0, f_y, c_y], [0.0, 0.0, 1.0]])
            intrinsic_matrices = k.unsqueeze(0).repeat(self.num_envs, 1, 1)
        self.intrinsic_matrices = intrinsic_matrices.to(self.device)
        starts, dirs = patterns.pinhole_camera_pattern(pc, self.intrinsic_matrices, "cpu")
        self.ray_starts, self.ray_directions = starts.to(self.device), dirs.to(self.device)   # [N, H*W, 3]
        self.num_rays = starts.shape[1]
        self.frame = torch.zeros(self.num_envs, dtype=torch.long, device=self.device)
        self.output = {}
        self.ray_hits_w = None

    def update(self, pos_w: torch.Tensor, quat_w: torch.Tensor) -> dict:
        cfg, N, R = self.cfg, self.num_envs, self.num_rays
        self.frame += 1
        q = quat_w.unsqueeze(1).expand(N, R, 4)
        starts_w = M.quat_apply(q, self.ray_starts) + pos_w.unsqueeze(1)
        dirs_w = M.quat_apply(q, self.ray_directions)
        want_dist = any(n in cfg.data_types for n in ("distance_to_image_plane", "distance_to_camera"))
        self.ray_hits_w, depth, normal, _ = raycast_mesh(starts_w, dirs_w, self.mesh, max_dist=1e6, return_distance=want_dist,
                                                         return_normal="normals" in cfg.data_types)
        H, W = self.image_shape
        if "distance_to_image_plane" in cfg.data_types:
            d = M.quat_apply(_quat_inv(quat_w).unsqueeze(1).expand(N, R, 4), depth.unsqueeze(-1) * dirs_w)[..., 0]
            if cfg.depth_clipping_behavior == "max":
                d = torch.clip(d, max=cfg.max_distance)
                d[torch.isnan(d)] = cfg.max_distance
            elif cfg.depth_clipping_behavior == "zero":
                d[d > cfg.max_distance] = 0.0
                d[torch.isnan(d)] = 0.0
            self.output["distance_to_image_plane"] = d.view(N, H, W, 1)
        if "distance_to_camera" in cfg.data_types:
            r = depth.clone()
            if cfg.depth_clipping_behavior == "max":
                r = torch.clip(r, max=cfg.max_distance)
            elif cfg.depth_clipping_behavior == "zero":
                r[r > cfg.max_distance] = 0.0
            self.output["distance_to_camera"] = r.view(N, H, W, 1)
        if "normals" in cfg.data_types:
            self.output["normals"] = normal.view(N, H, W, 3)
        return self.output
