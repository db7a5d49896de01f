"""Ray casting against a static triangle mesh on libb200lab (replaces the Warp path of the reference).

Mirrors LAB/utils/warp/ops.py: `convert_to_warp_mesh(points, indices, device)` -> mesh handle and
`raycast_mesh(ray_starts, ray_directions, mesh, max_dist, return_distance, return_normal, return_face_id)`
-> `(ray_hits, ray_distance | None, ray_normal | None, ray_face_id | None)` with `inf` on a miss
(ops.py:24-127), and `grid_pattern` of LAB/sensors/ray_caster/patterns/patterns.py:16-58.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .. import _lib


class WarpFreeMesh:
    """Device-side acceleration structure for one static mesh (what `wp.Mesh` is to the reference)."""

    def __init__(self, points, indices, cell_size: float = 0.0):
        self.points = np.ascontiguousarray(np.asarray(points, dtype=np.float32).reshape(-1, 3))
        self.indices = np.ascontiguousarray(np.asarray(indices, dtype=np.int32).reshape(-1, 3))
        L = _lib.lib()
        self._L = L
        self.handle = L.b200lab_mesh_create(self.points.ctypes.data, self.points.shape[0], self.indices.ctypes.data,
                                            self.indices.shape[0], float(cell_size))
        if not self.handle:
            raise RuntimeError("b200lab_mesh_create failed: " + L.b200lab_last_error().decode())
        self.device_ptr = L.b200lab_mesh_device_ptr(self.handle)
        self.nbytes = int(L.b200lab_mesh_bytes(self.handle))

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self._L.b200lab_mesh_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


def convert_to_warp_mesh(points, indices, device: str = "cuda") -> WarpFreeMesh:
    """Name kept from LAB/utils/warp/ops.py:130-145 so call sites do not change."""
    return WarpFreeMesh(points, indices)


def raycast_mesh(ray_starts: torch.Tensor, ray_directions: torch.Tensor, mesh: WarpFreeMesh, max_dist: float = 1e6,
                 return_distance: bool = False, return_normal: bool = False, return_face_id: bool = False):
    shape = ray_starts.shape
    dev = ray_starts.device
    starts = ray_starts.reshape(-1, 3).contiguous().float()
    dirs = ray_directions.reshape(-1, 3).contiguous().float()
    n = starts.shape[0]
    hits = torch.empty(n, 3, device=dev)
    dist = torch.empty(n, device=dev) if return_distance else None
    normal = torch.empty(n, 3, device=dev) if return_normal else None
    face = torch.empty(n, dtype=torch.int32, device=dev) if return_face_id else None
    p = _lib.ptr
    _lib.check(_lib.lib().b200lab_raycast_mesh(mesh.handle, p(starts), p(dirs), n, float(max_dist), p(hits), p(dist),
                                              p(normal), p(face), _lib.current_stream()))
    out_d = dist.view(shape[:-1]) if return_distance else None
    out_n = normal.view(shape) if return_normal else None
    out_f = face.view(shape[:-1]) if return_face_id else None
    return hits.view(shape), out_d, out_n, out_f


def grid_pattern(size, resolution: float, direction=(0.0, 0.0, -1.0), ordering: str = "xy"):
    """Ray starts / directions of GridPatternCfg as numpy [R,3] (patterns.py:16-58; torch.arange/meshgrid there)."""
    if ordering not in ("xy", "yx"):
        raise ValueError(f"Ordering must be 'xy' or 'yx'. Received: '{ordering}'.")
    if resolution <= 0:
        raise ValueError(f"Resolution must be greater than 0. Received: '{resolution}'.")
    x = torch.arange(start=-size[0] / 2, end=size[0] / 2 + 1.0e-9, step=resolution)
    y = torch.arange(start=-size[1] / 2, end=size[1] / 2 + 1.0e-9, step=resolution)
    gx, gy = torch.meshgrid(x, y, indexing="xy" if ordering == "xy" else "ij")
    starts = torch.zeros(gx.numel(), 3)
    starts[:, 0] = gx.flatten()
    starts[:, 1] = gy.flatten()
    dirs = torch.zeros_like(starts)
    dirs[..., :] = torch.tensor(list(direction))
    return starts.numpy(), dirs.numpy()
