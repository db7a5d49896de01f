"""GPU tests of the observation entry points added for pipelined consumers: any partition of the env range through
b200lab_compute_obs_range gives the rows b200lab_compute_obs gives; the speculative height scan is a no-op for sensor
configurations it does not cover; post_step's derived quantities are dropped by calls that advance the env."""
import copy

import pytest
import torch

from tests.mdp_common import init_fused_state, make_spec

pytestmark = pytest.mark.gpu


def _env(task, N, seed=5):
    from isaaclab_b200.envs.fused_env import FusedEnv
    from isaaclab_b200.sim.synthetic import SyntheticSim, generate_slabs, initial_env_state
    spec = make_spec(task, N)
    slabs = generate_slabs(spec, 4, seed=321)
    fus = FusedEnv(copy.deepcopy(spec), device="cuda", seed=seed)
    init = initial_env_state(spec, seed=9)
    init["episode_length_buf"][::6] = spec["max_episode_length"] - 1   # some envs reset in the first step
    init_fused_state(fus, init)
    sim = SyntheticSim(spec, K=4, device="cuda", slabs={k: v.clone() for k, v in slabs.items()})
    return spec, fus, sim


@pytest.mark.parametrize("task", ["Isaac-Velocity-Rough-Anymal-C-v0", "Isaac-Cartpole-v0"])
@pytest.mark.parametrize("ranges", [[(0, 100), (100, 101), (101, 256)], [(128, 256), (0, 128)]])
def test_compute_obs_range_partition_matches_full(task, ranges):
    N = 256
    g = torch.Generator().manual_seed(2)
    action = torch.randn(N, make_spec(task, N)["A"], generator=g).cuda()
    outs, resets = [], 0
    for mode in ("full", "ranges"):
        spec, fus, sim = _env(task, N)
        for k in range(2):
            views, prev = sim.views(k + 1), sim.views(k)[-1]
            if mode == "full":
                fus.step(action, views, prev)
            else:
                seen = []
                fus.step(action, views, prev, obs_ranges=ranges, after_range=lambda b, e: seen.append((b, e)))
                assert seen == ranges
            resets += int(fus.reset_buf.sum())
        torch.cuda.synchronize()
        outs.append((fus.obs_buf.clone(), fus.reward_buf.clone(), fus.reset_buf.clone()))
    assert resets > 0, "no env reset in the test horizon"
    for a, b, name in zip(outs[0], outs[1], ("obs", "reward", "reset_buf")):
        assert torch.equal(a, b), f"{task}: {name} differs between compute_obs and its range partition"


def test_compute_obs_range_rejects_bad_range():
    from isaaclab_b200 import _lib
    spec, fus, sim = _env("Isaac-Cartpole-v0", 64)
    v = sim.views(1)[-1]
    for b, e in ((-1, 4), (8, 4), (0, 65)):
        rc = fus.L.b200lab_compute_obs_range(fus.handle, v.ref, 0, b, e, _lib.current_stream())
        assert rc != 0 and b"range" in fus.L.b200lab_last_error()


def test_height_scan_prefetch_is_noop_without_height_scanner():
    """Cartpole has no ray caster: the call returns OK, launches nothing, and the step is unchanged."""
    spec, fus, sim = _env("Isaac-Cartpole-v0", 64)
    n0 = fus.L.b200lab_launch_count()
    fus.height_scan_prefetch(sim.views(1)[-1])
    assert fus.L.b200lab_launch_count() == n0


def test_derived_quantities_follow_state_written_between_steps():
    """compute_obs after post_step uses the quantities post_step derived; after a call that advances the env
    (here: env_update through set_uniforms) it derives them again from the tensors it is given."""
    N = 64
    spec, fus, sim = _env("Isaac-Velocity-Flat-Anymal-C-v0", N)
    action = torch.zeros(N, spec["A"], device="cuda")
    views, prev = sim.views(1), sim.views(0)[-1]
    fus.step(action, views, prev)
    obs_a = fus.compute_obs(views[-1]).clone()      # same state, derived quantities from post_step
    views[-1].root_state_w[:, 7:10] += 1.0           # the caller rewrites the base velocity ...
    fus.set_uniforms(None, None, None)               # ... and tells the library (b200lab_env_update)
    obs_b = fus.compute_obs(views[-1]).clone()
    torch.cuda.synchronize()
    assert not torch.equal(obs_a[:, :3], obs_b[:, :3]), "base_lin_vel columns did not follow the rewritten state"
    assert torch.equal(obs_a[:, 3:], obs_b[:, 3:]), "columns that do not depend on the base velocity changed"
