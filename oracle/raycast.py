"""Oracle for `raycast_mesh` (LAB/utils/warp/ops.py:24-127): brute-force fp64 closest hit (test infrastructure only).

Uses the C restatement (oracle/raycast_bruteforce.c, built by oracle/build.py) when it is built, else the same
algorithm in chunked numpy.  PARITY UNPINNED against warp-lang (see oracle/__init__.py)."""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

_LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_build", "liboracle_raycast.so")
_lib = None


def _clib():
    global _lib
    if _lib is None and os.path.exists(_LIB):
        _lib = ctypes.CDLL(_LIB)
        _lib.oracle_raycast_mesh.restype = None
        _lib.oracle_raycast_mesh.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.c_int64, ctypes.c_double, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.c_void_p]
        _lib.oracle_grid_build.restype = ctypes.c_void_p
        _lib.oracle_grid_build.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64, ctypes.c_float]
        _lib.oracle_grid_free.argtypes = [ctypes.c_void_p]
        _lib.oracle_raycast_mesh_grid.restype = None
        _lib.oracle_raycast_mesh_grid.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                                  ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_double,
                                                  ctypes.c_void_p]
    return _lib


def raycast_numpy(V, F, starts, dirs, max_dist, chunk=128):
    V = V.astype(np.float64)
    a, b, c = V[F[:, 0]], V[F[:, 1]], V[F[:, 2]]
    e1, e2 = b - a, c - a
    n = starts.shape[0]
    t_best = np.full(n, np.inf)
    f_best = np.full(n, -1, dtype=np.int32)
    for s in range(0, n, chunk):
        o = starts[s:s + chunk].astype(np.float64)[:, None, :]
        d = dirs[s:s + chunk].astype(np.float64)[:, None, :]
        p = np.cross(d, e2[None])
        det = (e1[None] * p).sum(-1)
        ok = np.abs(det) >= 1e-12
        inv = np.where(ok, 1.0 / np.where(ok, det, 1.0), 0.0)
        tv = o - a[None]
        u = (tv * p).sum(-1) * inv
        q = np.cross(tv, e1[None])
        v = (d * q).sum(-1) * inv
        t = (e2[None] * q).sum(-1) * inv
        hit = ok & (u >= 0) & (u <= 1) & (v >= 0) & (u + v <= 1) & (t > 0) & (t <= max_dist)
        t = np.where(hit, t, np.inf)
        k = t.argmin(axis=1)
        tb = t[np.arange(t.shape[0]), k]
        t_best[s:s + chunk] = tb
        f_best[s:s + chunk] = np.where(np.isfinite(tb), k, -1)
    return t_best, f_best


class Mesh:
    def __init__(self, vertices, faces):
        self.V = np.ascontiguousarray(np.asarray(vertices, dtype=np.float32).reshape(-1, 3))
        self.F = np.ascontiguousarray(np.asarray(faces, dtype=np.int32).reshape(-1, 3))

    def enable_grid(self, cell=0.25):
        """CPU-BASELINE mode (bench.py only): xy cell pre-filter for vertical rays; same results as brute force."""
        lib = _clib()
        if lib is None:
            raise RuntimeError("oracle C library not built (python oracle/build.py)")
        self._grid = lib.oracle_grid_build(self.V.ctypes.data, self.V.shape[0], self.F.ctypes.data, self.F.shape[0], cell)

    def raycast(self, starts, dirs, max_dist=1e6, return_t=False):
        if getattr(self, "_grid", None) and not return_t:
            s = np.ascontiguousarray(starts.detach().cpu().numpy().astype(np.float32))
            d = np.ascontiguousarray(dirs.detach().cpu().numpy().astype(np.float32))
            hits = np.empty((s.shape[0], 3), dtype=np.float32)
            _clib().oracle_raycast_mesh_grid(self._grid, self.V.ctypes.data, self.F.ctypes.data, self.F.shape[0],
                                             s.ctypes.data, d.ctypes.data, s.shape[0], float(max_dist), hits.ctypes.data)
            return torch.from_numpy(hits)
        """starts/dirs: torch [n,3] f32 -> hits torch [n,3] f32 (inf on miss)."""
        s = np.ascontiguousarray(starts.detach().cpu().numpy().astype(np.float32))
        d = np.ascontiguousarray(dirs.detach().cpu().numpy().astype(np.float32))
        n = s.shape[0]
        lib = _clib()
        if lib is not None:
            hits = np.empty((n, 3), dtype=np.float32)
            t = np.empty(n, dtype=np.float64)
            f = np.empty(n, dtype=np.int32)
            lib.oracle_raycast_mesh(self.V.ctypes.data, self.F.ctypes.data, self.F.shape[0], s.ctypes.data,
                                    d.ctypes.data, n, float(max_dist), hits.ctypes.data, t.ctypes.data, f.ctypes.data)
        else:
            t, f = raycast_numpy(self.V, self.F, s, d, max_dist)
            hits = (s.astype(np.float64) + t[:, None] * d.astype(np.float64)).astype(np.float32)
            hits[~np.isfinite(t)] = np.inf
        out = torch.from_numpy(hits)
        if return_t:
            return out, torch.from_numpy(t), torch.from_numpy(f)
        return out
