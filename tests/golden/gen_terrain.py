"""Golden fixtures of the terrain generators, produced by the UNMODIFIED reference functions
(LAB/terrains/height_field/{utils,hf_terrains}.py, LAB/terrains/trimesh/mesh_terrains.py) with ROUGH_TERRAINS_CFG's
sub-terrain parameters.  Run in the authoring container: python -m tests.golden.gen_terrain -> tests/golden/terrain.pt

* `convert_height_field_to_mesh` (pure numpy): vertices / triangles of a random int16 field with the slope-threshold rule;
* `pyramid_sloped_terrain` (normal / inverted) and `random_uniform_terrain` through the `height_field_to_mesh` decorator's
  own border logic: the assembled int16 height field, the origin (random tile: `np.random.seed(seed)`, replayed by the
  test through `np.random.RandomState(seed)`);
* mesh terrains (`pyramid_stairs_terrain`, `inverted_pyramid_stairs_terrain`, `random_grid_terrain`): trimesh is not in the
  image, so `trimesh.creation.box` / `translation_matrix` / `Trimesh` get a 20-line stand-in (an axis-aligned box IS its 8
  corners); what is pinned is the surface the reference's placement logic produces -- the height seen by vertical rays on a
  dense grid of points (oracle brute force), plus the origin; the random grid's torch draws are recorded."""
import copy
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from tests.golden import ref_harness  # noqa: E402

ref_harness.install()

_F = np.array([[0, 2, 1], [0, 3, 2], [4, 5, 6], [4, 6, 7], [0, 1, 5], [0, 5, 4], [1, 2, 6], [1, 6, 5], [2, 3, 7], [2, 7, 6],
               [3, 0, 4], [3, 4, 7]])


class _Trimesh:
    def __init__(self, vertices=None, faces=None):
        self.vertices, self.faces = np.asarray(vertices, dtype=np.float64), np.asarray(faces, dtype=np.int64)


def _box(extents, transform=None):
    ex = np.asarray(extents, dtype=np.float64) / 2
    c = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]], dtype=np.float64)
    v = c * ex
    if transform is not None:
        v = v + np.asarray(transform)[:3, 3]
    return _Trimesh(v, _F)


def _translation(pos):
    m = np.eye(4)
    m[:3, 3] = pos
    return m


def install_trimesh_stand_in():
    tm = types.ModuleType("trimesh")
    tm.Trimesh = _Trimesh
    tm.creation = types.SimpleNamespace(box=_box)
    tm.transformations = types.SimpleNamespace(translation_matrix=_translation)
    for mod in ("isaaclab.terrains.trimesh.mesh_terrains", "isaaclab.terrains.trimesh.utils", "isaaclab.terrains.height_field.utils"):
        m = sys.modules.get(mod)
        if m is not None:
            m.trimesh = tm
    return tm


def sample_heights(meshes, size, n=97):
    """Top surface seen by vertical rays on an n x n grid strictly inside the tile (points offset off every lattice)."""
    from oracle import raycast as rc
    V, F, off = [], [], 0
    for m in meshes:
        V.append(np.asarray(m.vertices, dtype=np.float32))
        F.append(np.asarray(m.faces, dtype=np.int64) + off)
        off += len(m.vertices)
    mesh = rc.Mesh(np.concatenate(V), np.concatenate(F).astype(np.int32))
    xs = (np.arange(n) + 0.5) / n * size[0] * 0.999 + 0.00037
    ys = (np.arange(n) + 0.5) / n * size[1] * 0.999 + 0.00053
    gx, gy = np.meshgrid(xs, ys, indexing="ij")
    s = torch.tensor(np.stack([gx.ravel(), gy.ravel(), np.full(gx.size, 50.0)], 1), dtype=torch.float32)
    d = torch.zeros_like(s)
    d[:, 2] = -1.0
    return mesh.raycast(s, d, 1e3)[:, 2].view(n, n).clone(), (xs, ys)


def main():
    import isaaclab.terrains.height_field.hf_terrains as hf
    import isaaclab.terrains.height_field.utils as hfu
    import isaaclab.terrains.trimesh.mesh_terrains as mt
    import isaaclab.terrains.trimesh.utils  # noqa: F401
    from isaaclab.terrains.config.rough import ROUGH_TERRAINS_CFG as REF_CFG
    install_trimesh_stand_in()
    out = {}
    # 1. convert_height_field_to_mesh
    rng = np.random.RandomState(3)
    field = (rng.randint(-30, 30, size=(23, 17)) * (rng.rand(23, 17) < 0.4)).astype(np.int16)
    v, t = hfu.convert_height_field_to_mesh(field, 0.1, 0.005, 0.75)
    v2, t2 = hfu.convert_height_field_to_mesh(field, 0.1, 0.005, None)
    out["convert"] = dict(field=torch.from_numpy(field), hs=0.1, vs=0.005, thr=0.75, vertices=torch.from_numpy(v.copy()),
                          triangles=torch.from_numpy(t.astype(np.int64)), vertices_nothr=torch.from_numpy(v2.copy()))
    # 2. height-field tiles through the reference decorator (the stand-in Trimesh keeps the vertices; the assembled int16
    #    field is recovered from vertex z / vertical_scale only where no slope correction moved x/y -> store the mesh)
    subs = REF_CFG.sub_terrains
    tiles = {}
    for name in ("hf_pyramid_slope", "hf_pyramid_slope_inv"):
        for dff in (0.0, 0.37, 1.0):
            cfg = copy.deepcopy(subs[name])
            cfg.size = (8.0, 8.0)
            meshes, origin = cfg.function(dff, cfg)
            tiles[(name, dff)] = dict(vertices=torch.from_numpy(np.asarray(meshes[0].vertices, dtype=np.float32)),
                                      faces=torch.from_numpy(np.asarray(meshes[0].faces, dtype=np.int64)),
                                      origin=torch.from_numpy(np.asarray(origin, dtype=np.float64)))
    for seed in (11, 12):
        cfg = copy.deepcopy(subs["random_rough"])
        cfg.size = (8.0, 8.0)
        np.random.seed(seed)
        meshes, origin = cfg.function(0.5, cfg)
        tiles[("random_rough", seed)] = dict(vertices=torch.from_numpy(np.asarray(meshes[0].vertices, dtype=np.float32)),
                                             faces=torch.from_numpy(np.asarray(meshes[0].faces, dtype=np.int64)),
                                             origin=torch.from_numpy(np.asarray(origin, dtype=np.float64)))
    first = None   # the triangle index array is the same for every tile of this size: keep one copy
    for t in tiles.values():
        f = t.pop("faces").to(torch.int32)
        assert first is None or torch.equal(first, f)
        first = f
    out["hf_faces"] = first
    out["hf_tiles"] = tiles
    # 3. mesh tiles: sampled top surface
    mesh_tiles = {}
    for name in ("pyramid_stairs", "pyramid_stairs_inv"):
        for dff in (0.0, 0.37, 1.0):
            cfg = copy.deepcopy(subs[name])
            cfg.size = (8.0, 8.0)
            meshes, origin = cfg.function(dff, cfg)
            z, _ = sample_heights(meshes, cfg.size)
            mesh_tiles[(name, dff)] = dict(heights=z, origin=torch.from_numpy(np.asarray(origin, dtype=np.float64)))
    for seed in (21, 22):
        cfg = copy.deepcopy(subs["boxes"])
        cfg.size = (8.0, 8.0)
        torch.manual_seed(seed)
        draws = []
        orig = torch.Tensor.uniform_

        def rec(self, a=0.0, b=1.0, **kw):
            r = orig(self, a, b, **kw)
            draws.append(r.clone())
            return r

        torch.Tensor.uniform_ = rec
        try:
            meshes, origin = cfg.function(0.6, cfg)
        finally:
            torch.Tensor.uniform_ = orig
        z, _ = sample_heights(meshes, cfg.size)
        mesh_tiles[("boxes", seed)] = dict(heights=z, origin=torch.from_numpy(np.asarray(origin, dtype=np.float64)),
                                           draws=draws[0].double())
    out["mesh_tiles"] = mesh_tiles
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "terrain.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB", list(tiles)[:2], list(mesh_tiles)[:2])


if __name__ == "__main__":
    main()
