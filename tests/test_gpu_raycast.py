"""GPU tests of the stand-alone `b200lab_raycast_mesh` entry point (isaaclab_b200.sensors.ray_caster.raycast_mesh, the
replacement of LAB/utils/warp/ops.py:24-127) against the fp64 brute-force oracle: vertical, tilted and horizontal rays
through the cell walk, `max_dist` cut, miss = inf, `ray_distance`, `ray_normal`, `ray_face_id`.

Tolerance: 1e-4 m on hit distance / position (north_star).  PARITY UNPINNED against warp-lang's mesh_query_ray (absent
from the image); the contract is the one kernels.py:13-75 documents."""
import numpy as np
import pytest
import torch

from oracle import build as obuild
from oracle import raycast as rc
from tests.mdp_common import small_terrain

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _built():
    obuild.build()


@pytest.fixture(scope="module")
def terrain_mesh():
    from isaaclab_b200.sensors.ray_caster import convert_to_warp_mesh
    terr = small_terrain()
    V, F = terr.terrain_mesh
    V, F = np.asarray(V, dtype=np.float32), np.asarray(F, dtype=np.int32)
    return V, F, rc.Mesh(V, F), convert_to_warp_mesh(V, F)


def _rays(V, n, seed, kind):
    g = torch.Generator().manual_seed(seed)
    lo, hi = torch.tensor(V.min(0)), torch.tensor(V.max(0))
    s = torch.rand(n, 3, generator=g) * (hi - lo) + lo
    d = torch.zeros(n, 3)
    if kind == "vertical":
        s[:, 2] = hi[2] + 5.0
        d[:, 2] = -1.0
    elif kind == "up":           # from below the terrain, pointing up (two-sided triangles)
        s[:, 2] = lo[2] - 3.0
        d[:, 2] = 1.0
    elif kind == "tilted":
        s[:, 2] = hi[2] + 2.0
        d[:, 2] = -1.0
        d[:, :2] = 0.6 * torch.randn(n, 2, generator=g)
    elif kind == "grazing":      # nearly horizontal, starting inside the height range of the rough tiles
        s[:, 2] = lo[2] + torch.rand(n, generator=g) * (hi[2] - lo[2]) + 0.05
        ang = torch.rand(n, generator=g) * 2 * np.pi
        d[:, 0], d[:, 1] = torch.cos(ang), torch.sin(ang)
        d[:, 2] = -0.05 * torch.rand(n, generator=g)
    elif kind == "horizontal":   # d.z == 0 exactly
        s[:, 2] = lo[2] + torch.rand(n, generator=g) * (hi[2] - lo[2])
        ang = torch.rand(n, generator=g) * 2 * np.pi
        d[:, 0], d[:, 1] = torch.cos(ang), torch.sin(ang)
    elif kind == "axis":         # rays along +x / -y exactly (one DDA axis degenerate)
        s[:, 2] = lo[2] + torch.rand(n, generator=g) * (hi[2] - lo[2])
        d[: n // 2, 0] = 1.0
        d[n // 2:, 1] = -1.0
    d = d / d.norm(dim=-1, keepdim=True)
    return s, d


def _check(oracle, gpu_mesh, s, d, max_dist, V, F, tol=1e-4):
    from isaaclab_b200.sensors.ray_caster import raycast_mesh
    h_o, t_o, f_o = oracle.raycast(s, d, max_dist, return_t=True)
    h, dist, nrm, face = raycast_mesh(s.cuda(), d.cuda(), gpu_mesh, max_dist, True, True, True)
    torch.cuda.synchronize()
    h, dist, nrm, face = h.cpu(), dist.cpu(), nrm.cpu(), face.cpu()
    hit_o = torch.isfinite(t_o)
    hit = torch.isfinite(dist)
    # rays whose oracle hit lies within a hair of max_dist may legitimately flip
    flips = hit != hit_o
    near_cut = (t_o - max_dist).abs() < 1e-3
    assert int((flips & ~near_cut).sum()) == 0, f"{int((flips & ~near_cut).sum())} hit/miss mismatches"
    both = hit & hit_o
    # closest-hit distance: 1e-4 m.  A grazing ray that passes within fp32 rounding of a shared edge may land on the
    # neighbouring surface: bounded count, everything else within tol.
    err = (dist[both].double() - t_o[both]).abs()
    bad = err > tol
    assert float(bad.float().mean()) < 2e-3, f"{int(bad.sum())} of {int(both.sum())} hit distances off by > {tol}"
    assert float(err[~bad].max()) <= tol
    # miss = inf everywhere (ops.py:70), face id -1
    assert torch.isinf(h[~hit]).all() and torch.isinf(dist[~hit]).all() and (face[~hit] == -1).all()
    assert torch.isinf(nrm[~hit]).all()
    # hit = start + t * dir (kernels.py:68)
    torch.testing.assert_close(h[hit], (s + dist.unsqueeze(-1) * d)[hit], rtol=1e-5, atol=1e-4)
    # face id indexes the caller's face array; the hit lies in that face's plane and inside it; normal = unit face normal
    ok = both.clone()
    ok[both] = ~bad
    fid = face[ok].long()
    assert (fid >= 0).all() and (fid < F.shape[0]).all()
    tri = torch.from_numpy(V[F[fid.numpy()]]).double()        # [k,3,3]
    n_ref = torch.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0], dim=-1)
    n_ref = n_ref / n_ref.norm(dim=-1, keepdim=True)
    plane = ((h[ok].double() - tri[:, 0]) * n_ref).sum(-1).abs()
    assert float(plane.max()) < 2e-4, "hit point is not on the reported face"
    cosang = (nrm[ok].double() * n_ref).sum(-1).abs()
    assert float(cosang.min()) > 1 - 1e-5, "ray_normal is not the face normal"
    # The face id itself may differ from the oracle's where faces coincide (coplanar overlapping patches of the terrain
    # border / tile floors) or share the edge hit: require the hit to be INSIDE the reported face instead.
    e1, e2, w = tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0], h[ok].double() - tri[:, 0]
    d11, d12, d22 = (e1 * e1).sum(-1), (e1 * e2).sum(-1), (e2 * e2).sum(-1)
    w1, w2 = (w * e1).sum(-1), (w * e2).sum(-1)
    den = d11 * d22 - d12 * d12
    bu, bv = (d22 * w1 - d12 * w2) / den, (d11 * w2 - d12 * w1) / den
    slack = 2e-4 / torch.sqrt(torch.minimum(d11, d22))     # 0.2 mm on the shorter edge
    inside = (bu >= -slack) & (bv >= -slack) & (bu + bv <= 1 + slack)
    assert bool(inside.all()), f"{int((~inside).sum())} hit points lie outside the reported face"
    return int(both.sum()), int(hit_o.numel())


@pytest.mark.parametrize("kind", ["vertical", "up", "tilted", "grazing", "horizontal", "axis"])
def test_raycast_mesh_matches_bruteforce(terrain_mesh, kind):
    V, F, oracle, gpu = terrain_mesh
    s, d = _rays(V, 4096, seed=11, kind=kind)
    n_hit, n = _check(oracle, gpu, s, d, 1e6, V, F)
    assert n_hit > (n // 2 if kind in ("vertical", "up", "tilted") else 50), f"{kind}: only {n_hit} of {n} rays hit"


def test_raycast_mesh_max_dist_and_misses(terrain_mesh):
    from isaaclab_b200.sensors.ray_caster import raycast_mesh
    V, F, oracle, gpu = terrain_mesh
    s, d = _rays(V, 2048, seed=12, kind="tilted")
    _, t_o, _ = oracle.raycast(s, d, 1e6, return_t=True)
    cut = float(t_o[torch.isfinite(t_o)].median())      # roughly half of the rays miss under this cut
    n_hit, n = _check(oracle, gpu, s, d, cut, V, F)
    assert 0.2 * n < n_hit < 0.8 * n
    s2 = s.clone()
    s2[:, 0] += 1e4                                      # rays that never enter the mesh footprint
    h, dist, _, face = raycast_mesh(s2.cuda(), d.cuda(), gpu, 1e6, True, False, True)
    assert torch.isinf(h).all() and torch.isinf(dist).all() and (face == -1).all()


def test_raycast_mesh_optional_outputs_and_empty_batch(terrain_mesh):
    from isaaclab_b200.sensors.ray_caster import raycast_mesh
    V, F, oracle, gpu = terrain_mesh
    s, d = _rays(V, 512, seed=13, kind="vertical")
    h, dist, nrm, face = raycast_mesh(s.cuda(), d.cuda(), gpu)
    assert dist is None and nrm is None and face is None and h.shape == (512, 3)
    h_o = oracle.raycast(s, d, 1e6)
    fin = torch.isfinite(h_o[:, 2])
    assert torch.equal(fin, torch.isfinite(h[:, 2]).cpu())
    torch.testing.assert_close(h.cpu()[fin], h_o[fin], rtol=0, atol=1e-4)
    # batched shapes [N, R, 3] as RayCaster passes them (ray_caster.py:255)
    hb, db, _, _ = raycast_mesh(s.cuda().view(8, 64, 3), d.cuda().view(8, 64, 3), gpu, return_distance=True)
    assert hb.shape == (8, 64, 3) and db.shape == (8, 64)
    assert torch.equal(hb.view(-1, 3), h)
    he, _, _, _ = raycast_mesh(torch.zeros(0, 3, device="cuda"), torch.zeros(0, 3, device="cuda"), gpu)
    assert he.shape == (0, 3)


def test_raycast_mesh_known_answers():
    """Flat square at z = 0.5 under one at z = 1.5: the CPU known-answer cases of tests/test_oracle_raycast.py on the GPU."""
    from isaaclab_b200.sensors.ray_caster import convert_to_warp_mesh, raycast_mesh
    V = np.array([[0, 0, .5], [2, 0, .5], [2, 2, .5], [0, 2, .5], [0, 0, 1.5], [2, 0, 1.5], [2, 2, 1.5], [0, 2, 1.5]],
                 dtype=np.float32)
    F = np.array([[0, 1, 2], [0, 2, 3], [4, 5, 6], [4, 6, 7]], dtype=np.int32)
    m = convert_to_warp_mesh(V, F)
    s = torch.tensor([[1.0, 0.5, 3.0], [1.0, 0.5, 1.0], [5.0, 5.0, 3.0], [1.0, 0.5, 0.2], [1.0, 0.5, 3.0], [-1.0, 1.0, 1.0]])
    d = torch.tensor([[0, 0, -1.0], [0, 0, -1.0], [0, 0, -1.0], [0, 0, -1.0], [0, 0, 1.0], [1.0, 0.0, 0.0]])
    h, dist, nrm, face = raycast_mesh(s.cuda(), d.cuda(), m, 100.0, True, True, True)
    h, dist, face = h.cpu(), dist.cpu(), face.cpu()
    assert abs(h[0, 2].item() - 1.5) < 1e-6 and abs(dist[0].item() - 1.5) < 1e-6 and face[0].item() in (2, 3)
    assert abs(h[1, 2].item() - 0.5) < 1e-6 and face[1].item() in (0, 1)
    assert torch.isinf(h[2]).all() and torch.isinf(h[3]).all() and torch.isinf(h[4]).all()
    assert torch.isinf(h[5]).all(), "a horizontal ray between the two layers hits neither"
    assert abs(abs(nrm.cpu()[0, 2].item()) - 1.0) < 1e-6


def test_ray_caster_camera_depth_images_match_brute_force():
    """LAB/sensors/ray_caster/ray_caster_camera.py:257-318 on the mesh kernel: tilted pinhole cameras over the terrain; depth
    along the optical axis, ray length and normals against the oracle's brute-force query pushed through the same formulas."""
    from isaaclab_b200.sensors import patterns as P
    from isaaclab_b200.sensors.ray_caster import WarpFreeMesh
    from isaaclab_b200.sensors.ray_caster_camera import RayCasterCamera, RayCasterCameraCfg
    from isaaclab_b200.utils import math as M
    from oracle import raycast as orc
    from tests.mdp_common import small_terrain
    V, F = small_terrain().terrain_mesh
    mesh = WarpFreeMesh(V, F)
    N = 6
    g = torch.Generator().manual_seed(4)
    cfg = RayCasterCameraCfg(pattern_cfg=P.PinholeCameraPatternCfg(width=16, height=12, focal_length=24.0, horizontal_aperture=36.0),
                             data_types=["distance_to_image_plane", "distance_to_camera", "normals"], max_distance=6.0,
                             depth_clipping_behavior="max")
    cam = RayCasterCamera(cfg, mesh, N)
    pos = torch.stack([torch.rand(N, generator=g) * 6 - 3, torch.rand(N, generator=g) * 6 - 3, 2.0 + torch.rand(N, generator=g)], dim=1)
    # cameras looking forward and down by 30-70 degrees, random heading (rotation about y, then about z)
    pitch, yaw = 0.5 + 0.7 * torch.rand(N, generator=g), 6.28 * torch.rand(N, generator=g)
    qy = torch.stack([torch.cos(pitch / 2), torch.zeros(N), torch.sin(pitch / 2), torch.zeros(N)], dim=1)
    qz = torch.stack([torch.cos(yaw / 2), torch.zeros(N), torch.zeros(N), torch.sin(yaw / 2)], dim=1)

    def qmul(a, b):
        w1, x1, y1, z1 = a.unbind(-1)
        w2, x2, y2, z2 = b.unbind(-1)
        return torch.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                            w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], dim=-1)

    quat = qmul(qz, qy)
    out = cam.update(pos.cuda(), quat.cuda())
    assert out["distance_to_image_plane"].shape == (N, 12, 16, 1) and out["normals"].shape == (N, 12, 16, 3)
    # the same images from the brute-force closest hit
    R = cam.num_rays
    q = quat.unsqueeze(1).expand(N, R, 4)
    starts = M.quat_apply(q, cam.ray_starts.cpu()) + pos.unsqueeze(1)
    dirs = M.quat_apply(q, cam.ray_directions.cpu())
    _, t, _ = orc.Mesh(V, F).raycast(starts.reshape(-1, 3), dirs.reshape(-1, 3), max_dist=1e6, return_t=True)
    t = t.float().view(N, R)
    assert torch.isfinite(t).float().mean() > 0.5
    want_cam = torch.clip(t, max=6.0)
    q_inv = torch.cat([quat[:, :1], -quat[:, 1:]], dim=1).unsqueeze(1).expand(N, R, 4)
    want_plane = torch.clip(M.quat_apply(q_inv, t.unsqueeze(-1) * dirs)[..., 0], max=6.0)
    want_plane[torch.isnan(want_plane)] = 6.0
    torch.testing.assert_close(out["distance_to_camera"].cpu().view(N, R), want_cam, rtol=1e-4, atol=2e-4)
    torch.testing.assert_close(out["distance_to_image_plane"].cpu().view(N, R), want_plane, rtol=1e-4, atol=2e-4)
    hit = torch.isfinite(t)
    nrm = out["normals"].cpu().view(N, R, 3)
    assert torch.allclose(nrm[hit].norm(dim=-1), torch.ones(int(hit.sum())), atol=1e-4)
    with pytest.raises(ValueError, match="does not support"):
        RayCasterCamera(RayCasterCameraCfg(pattern_cfg=cfg.pattern_cfg, data_types=["rgb"]), mesh, N)
