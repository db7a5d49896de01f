"""The shared-memory-staged height scan (compute_obs_patch_kernel, one CTA per env) against the per-ray kernel it replaces
on the hot path: same candidates, same arithmetic => BITWISE the same observation rows and ray hits, for every grid
geometry -- including the fallbacks (footprint spanning more grid rows than the staging area holds; entry / record
capacity overflow; large triangles that are never staged).  `B200LAB_OBS_IMPL` (read at env creation) selects the kernel.

Parity of either kernel with the oracle is tests/test_gpu_mdp.py / test_gpu_full_config.py; this file pins the two
implementations to each other so that the oracle-checked per-ray kernel vouches for the staged one at any size."""
import copy
import os

import pytest
import torch

from tests.mdp_common import make_full_spec, make_spec

pytestmark = pytest.mark.gpu


def _env(spec, impl, **kw):
    from isaaclab_b200.envs.fused_env import FusedEnv
    old = os.environ.get("B200LAB_OBS_IMPL")
    os.environ["B200LAB_OBS_IMPL"] = str(impl)
    try:
        return FusedEnv(copy.deepcopy(spec), device="cuda", seed=3, **kw)
    finally:
        if old is None:
            os.environ.pop("B200LAB_OBS_IMPL", None)
        else:
            os.environ["B200LAB_OBS_IMPL"] = old


def _compare(spec, steps=2, drift=0.3):
    from isaaclab_b200.sim.synthetic import SyntheticSim, generate_slabs
    slabs = generate_slabs(spec, steps, seed=21)
    g = torch.Generator().manual_seed(4)
    N = spec["N"]
    envs = [_env(spec, 0), _env(spec, 1), _env(spec, 1, store_ray_hits=False), _env(spec, 2), _env(spec, 3)]   # 2, 3: Q = 4, 2 columns / thread
    sims = [SyntheticSim(spec, K=steps, device="cuda", slabs={k: v.clone() for k, v in slabs.items()}) for _ in envs]
    u_obs = torch.rand(N, envs[0].obs_dim, generator=g).cuda()
    d = ((torch.rand(N, 3, generator=g) * 2 - 1) * drift).cuda()
    for k in range(steps):
        outs = []
        for env, sim in zip(envs, sims):
            env.set_uniforms(None, None, u_obs)
            env.rc_drift.copy_(d)
            env.rc_is_outdated.fill_(1)
            outs.append(env.compute_obs(sim.views(k)[-1]).clone())
        torch.cuda.synchronize()
        assert torch.isfinite(outs[0]).all()
        assert torch.equal(outs[0], outs[1]), f"step {k}: staged kernel differs from the per-ray kernel"
        assert torch.equal(outs[0], outs[2]), f"step {k}: observations change when the hit buffer is not stored"
        assert torch.equal(envs[0].rc_ray_hits_w, envs[1].rc_ray_hits_w), f"step {k}: ray hits differ"
        assert envs[2].rc_ray_hits_w is None
        for i in (3, 4):
            assert torch.equal(outs[0], outs[i]), f"step {k}: the {4 if i == 3 else 2}-column kernel differs from the per-ray kernel"
            assert torch.equal(envs[0].rc_ray_hits_w, envs[i].rc_ray_hits_w), f"step {k}: ray hits differ (multi-column kernel)"
        # perf mode: Philox noise, one counter block per thread in the multi-column kernels -- same stream, same bits
        philox = []
        for env, sim in zip(envs, sims):
            env.set_uniforms(None, None, None)
            env.rc_is_outdated.fill_(1)
            philox.append(env.compute_obs(sim.views(k)[-1]).clone())
        assert not torch.equal(philox[0], outs[0])
        for i in (1, 3, 4):
            assert torch.equal(philox[0], philox[i]), f"step {k}: Philox noise differs in kernel variant {i}"
    # a sensor that is NOT outdated re-reads its stored hits in both kernels
    for env, sim in zip(envs[:2], sims[:2]):
        env.rc_is_outdated.fill_(0)
        env.rc_ray_hits_w.mul_(0.5)
        outs.append(env.compute_obs(sim.views(0)[-1]).clone())
    assert torch.equal(outs[-1], outs[-2])
    return outs[0]


@pytest.mark.parametrize("task", ["Isaac-Velocity-Rough-Anymal-C-v0", "Isaac-Velocity-Rough-H1-v0"])
@pytest.mark.parametrize("cell", [0.0, 0.02, 0.033, 0.11, 0.6], ids=["auto", "tiny-cells", "small-cells", "coarse", "huge-cells"])
def test_staged_scan_is_bitwise_the_per_ray_scan(task, cell):
    spec = make_spec(task, 160)
    spec["raycaster"]["cell_size"] = cell
    _compare(spec)


def test_staged_scan_full_terrain_4096():
    obs = _compare(make_full_spec("Isaac-Velocity-Rough-Anymal-C-v0", 4096), steps=1)
    assert obs.shape == (4096, 235)


def test_store_ray_hits_false_needs_an_always_outdated_sensor():
    from isaaclab_b200.envs.fused_env import FusedEnv
    spec = make_spec("Isaac-Velocity-Rough-Anymal-C-v0", 32)
    spec["raycaster"]["update_period"] = 0.1
    with pytest.raises(ValueError, match="store_ray_hits"):
        FusedEnv(spec, device="cuda", store_ray_hits=False)
