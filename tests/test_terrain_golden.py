"""isaaclab_b200.terrains against fixtures produced by the UNMODIFIED reference generators (tests/golden/gen_terrain.py):
height-field -> mesh conversion incl. the slope-threshold rule (vertices bit-exact, same triangle set), the pyramid-slope
and random-uniform tiles through the border logic (bit-exact vertices given the same numpy stream), and the surface of the
mesh tiles (stairs, inverted stairs, random grid) as seen by vertical rays on a dense grid, plus every tile origin.

Not pinned: the order in which `TerrainGenerator` consumes its random stream across tiles (difficulty jitter, tile
choice) -- the reference mixes numpy's global state, a Generator and torch's RNG there; the per-tile functions are what
determines the geometry the ray caster sees."""
import copy
import os

import numpy as np
import pytest
import torch

from isaaclab_b200 import terrains as T
from oracle import build as obuild
from oracle import raycast as rc

FIX = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "terrain.pt")


@pytest.fixture(scope="module")
def fx():
    obuild.build()
    return torch.load(FIX, weights_only=False)


def _tri_set(f):
    return {tuple(r) for r in np.asarray(f, dtype=np.int64).tolist()}


def test_height_field_to_mesh_matches_reference(fx):
    c = fx["convert"]
    v, f = T.height_field_to_mesh(c["field"].numpy(), c["hs"], c["vs"], c["thr"])
    assert np.array_equal(v.astype(np.float32), c["vertices"].numpy()), "vertices (slope-threshold rule) differ"
    assert _tri_set(f) == _tri_set(c["triangles"].numpy())
    v2, _ = T.height_field_to_mesh(c["field"].numpy(), c["hs"], c["vs"], None)
    assert np.array_equal(v2.astype(np.float32), c["vertices_nothr"].numpy())
    assert not np.array_equal(v, v2), "the fixture does not exercise the slope correction"


def _sub(name):
    cfg = copy.deepcopy(T.ROUGH_TERRAINS_CFG.sub_terrains[name])
    cfg.size = (8.0, 8.0)
    return cfg


@pytest.mark.parametrize("name", ["hf_pyramid_slope", "hf_pyramid_slope_inv"])
def test_pyramid_sloped_tiles_match_reference(fx, name):
    for (n, dff), ref in fx["hf_tiles"].items():
        if n != name:
            continue
        v, f, origin = T.pyramid_sloped(dff, _sub(name), np.random.RandomState(0))
        assert np.array_equal(v.astype(np.float32), ref["vertices"].numpy()), f"{name} difficulty {dff}: vertices differ"
        assert _tri_set(f) == _tri_set(fx["hf_faces"].numpy())
        np.testing.assert_allclose(origin, ref["origin"].numpy(), rtol=0, atol=1e-12)


def test_random_uniform_tile_matches_reference_on_the_same_numpy_stream(fx):
    seen = 0
    for (n, seed), ref in fx["hf_tiles"].items():
        if n != "random_rough":
            continue
        v, f, origin = T.random_uniform(0.5, _sub("random_rough"), np.random.RandomState(seed))
        assert np.array_equal(v.astype(np.float32), ref["vertices"].numpy()), f"seed {seed}: vertices differ"
        np.testing.assert_allclose(origin, ref["origin"].numpy(), rtol=0, atol=1e-12)
        seen += 1
    assert seen == 2


def _heights(v, f, size, n=97):
    mesh = rc.Mesh(v.astype(np.float32), f.astype(np.int32))
    xs = (np.arange(n) + 0.5) / n * size[0] * 0.999 + 0.00037
    ys = (np.arange(n) + 0.5) / n * size[1] * 0.999 + 0.00053
    gx, gy = np.meshgrid(xs, ys, indexing="ij")
    s = torch.tensor(np.stack([gx.ravel(), gy.ravel(), np.full(gx.size, 50.0)], 1), dtype=torch.float32)
    d = torch.zeros_like(s)
    d[:, 2] = -1.0
    return mesh.raycast(s, d, 1e3)[:, 2].view(n, n)


class _Replay:
    """Feeds the recorded draws of the reference (torch `uniform_`, one per grid cell in x-major order)."""

    def __init__(self, draws):
        self.it = iter(draws.tolist())

    def uniform(self, lo, hi):
        return next(self.it)


@pytest.mark.parametrize("name", ["pyramid_stairs", "pyramid_stairs_inv", "boxes"])
def test_mesh_tiles_have_the_reference_surface(fx, name):
    seen = 0
    for (n, key), ref in fx["mesh_tiles"].items():
        if n != name:
            continue
        cfg = _sub(name)
        if name == "boxes":
            v, f, origin = T.random_grid(0.6, cfg, _Replay(ref["draws"]))
        else:
            v, f, origin = T.pyramid_stairs(key, cfg, np.random.RandomState(0))
        z = _heights(v, f, cfg.size)
        want = ref["heights"]
        assert torch.isfinite(z).all() and torch.isfinite(want).all()
        bad = (z - want).abs() > 1e-5
        assert int(bad.sum()) == 0, f"{name} {key}: {int(bad.sum())} of {bad.numel()} sampled heights differ (max {float((z - want).abs().max())})"
        np.testing.assert_allclose(origin, ref["origin"].numpy(), rtol=0, atol=1e-9)
        seen += 1
    assert seen >= 2
