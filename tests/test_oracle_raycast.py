"""Oracle ray query (oracle/raycast.py): known answers, C brute force == numpy brute force, and the grid-accelerated
CPU-baseline variant == brute force.  PARITY UNPINNED against warp-lang's mesh_query_ray (not in the image); the
contract checked here is the one LAB/utils/warp/ops.py:24-127 documents: closest hit within max_dist, inf on miss."""
import numpy as np
import pytest
import torch

from oracle import build as obuild
from oracle import raycast as rc
from tests.mdp_common import small_terrain


@pytest.fixture(scope="module", autouse=True)
def _built():
    obuild.build()


def _square(z=0.5):
    V = np.array([[0, 0, z], [2, 0, z], [2, 2, z], [0, 2, z]], dtype=np.float32)
    F = np.array([[0, 1, 2], [0, 2, 3]], dtype=np.int32)
    return rc.Mesh(V, F)


def test_known_answers_on_a_flat_square():
    m = _square(0.5)
    s = torch.tensor([[1.0, 1.0, 3.0], [0.5, 1.5, 3.0], [5.0, 5.0, 3.0], [1.0, 1.0, -1.0], [1.0, 1.0, 3.0]])
    d = torch.tensor([[0, 0, -1.0], [0, 0, -1.0], [0, 0, -1.0], [0, 0, -1.0], [0, 0, 1.0]])
    h = m.raycast(s, d, 100.0)
    assert torch.allclose(h[0], torch.tensor([1.0, 1.0, 0.5]))
    assert torch.allclose(h[1], torch.tensor([0.5, 1.5, 0.5]))
    assert torch.isinf(h[2]).all(), "outside the mesh -> inf"
    assert torch.isinf(h[3]).all(), "mesh behind the ray origin -> inf"
    assert torch.isinf(h[4]).all(), "pointing away -> inf"


def test_max_dist_cuts_hits():
    m = _square(0.0)
    s, d = torch.tensor([[1.0, 1.0, 3.0]]), torch.tensor([[0.0, 0.0, -1.0]])
    assert torch.isfinite(m.raycast(s, d, 3.5)).all()
    assert torch.isinf(m.raycast(s, d, 2.5)).all()


def test_closest_of_stacked_layers_wins():
    V = np.concatenate([_square(0.0).V, _square(1.0).V])
    F = np.concatenate([_square().F, _square().F + 4])
    m = rc.Mesh(V, F)
    h = m.raycast(torch.tensor([[1.0, 0.5, 3.0]]), torch.tensor([[0.0, 0.0, -1.0]]), 10.0)
    assert abs(h[0, 2].item() - 1.0) < 1e-6
    h = m.raycast(torch.tensor([[1.0, 0.5, 0.5]]), torch.tensor([[0.0, 0.0, -1.0]]), 10.0)
    assert abs(h[0, 2].item() - 0.0) < 1e-6


def test_empty_ray_batch():
    h = _square().raycast(torch.zeros(0, 3), torch.zeros(0, 3), 10.0)
    assert h.shape == (0, 3)


def _terrain_rays(n, seed, tilt=0.0):
    terr = small_terrain()
    V, F = np.asarray(terr.vertices, dtype=np.float32), np.asarray(terr.faces, dtype=np.int32)
    g = torch.Generator().manual_seed(seed)
    lo, hi = torch.tensor(V.min(0)), torch.tensor(V.max(0))
    s = torch.rand(n, 3, generator=g) * (hi - lo) + lo
    s[:, 2] = hi[2] + 5.0
    d = torch.zeros(n, 3)
    d[:, 2] = -1.0
    if tilt:
        d[:, :2] = tilt * torch.randn(n, 2, generator=g)
        d = d / d.norm(dim=-1, keepdim=True)
    return rc.Mesh(V, F), s, d


def test_c_bruteforce_matches_numpy_bruteforce():
    m, s, d = _terrain_rays(192, seed=4, tilt=0.2)
    h_c, t_c, f_c = m.raycast(s, d, 100.0, return_t=True)
    t_np, f_np = rc.raycast_numpy(m.V, m.F, s.numpy(), d.numpy(), 100.0)
    hit = np.isfinite(t_np)
    assert hit.sum() > 100
    assert np.array_equal(hit, np.isfinite(t_c.numpy()))
    np.testing.assert_allclose(t_c.numpy()[hit], t_np[hit], rtol=1e-9, atol=1e-9)


def test_grid_variant_matches_bruteforce_on_vertical_rays():
    """The cell pre-filter used as bench.py's CPU baseline returns what the brute force returns."""
    m, s, d = _terrain_rays(20000, seed=5)
    brute = m.raycast(s, d, 100.0)
    m.enable_grid()
    grid = m.raycast(s, d, 100.0)
    fin = torch.isfinite(brute[:, 2])
    assert torch.equal(fin, torch.isfinite(grid[:, 2]))
    assert fin.float().mean() > 0.95
    torch.testing.assert_close(grid[fin], brute[fin], rtol=0, atol=1e-6)
