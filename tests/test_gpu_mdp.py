"""GPU parity of the fused non-physics step (libb200lab through the C ABI) against the CPU oracle on the same seeded
synthetic simulator state and the same pre-drawn uniforms, for the five BASELINE configs' task definitions.

Bars (BASELINE.json north_star): reset/termination masks, episode counters, command-resample indices bit-exact;
float observations / rewards rtol 1e-5, atol 1e-6; height-scan hit distances 1e-4 m."""
import copy

import pytest
import torch

from oracle.mdp_env import OracleEnv
from tests.mdp_common import (close, draw_uniforms, exact, init_fused_state, init_oracle_state, make_spec,
                              run_step_parity, slab_views_cpu)

pytestmark = pytest.mark.gpu

TASKS = ["Isaac-Cartpole-v0", "Isaac-Velocity-Flat-Anymal-C-v0", "Isaac-Velocity-Rough-Anymal-C-v0",
         "Isaac-Velocity-Rough-Unitree-Go1-v0", "Isaac-Velocity-Rough-H1-v0",
         "B200-Opcode-Coverage-v0"]   # synthetic cfg: the opcodes / actuator kinds / options the five tasks do not use


# tasks without a reference fixture of their own (their cfgs are pinned by tests/test_task_cfgs_golden.py): CUDA vs oracle
MORE_TASKS = ["Isaac-Velocity-Rough-G1-v0", "Isaac-Velocity-Flat-Cassie-v0", "Isaac-Velocity-Rough-Unitree-A1-v0",
              "Isaac-Velocity-Rough-Anymal-D-v0"]


@pytest.mark.parametrize("task", MORE_TASKS)
def test_fused_step_matches_oracle_more_robots(task):
    """37-joint G1 (three implicit actuator groups, finger joints, joint-subset torque / acceleration penalties), Cassie,
    A1 (DC motor) and Anymal-D through the same kernels."""
    N, steps = (128, 5) if "Rough" in task else (256, 6)
    spec = make_spec(task, N)
    _, _, n_resets, _ = run_step_parity(spec, steps, fused=True, tag=task)
    assert n_resets > 0


@pytest.mark.parametrize("fused", [True, False, "spec"], ids=["fused-decimation", "per-substep", "fused+speculative-scan"])
@pytest.mark.parametrize("task", TASKS)
def test_fused_step_matches_oracle(task, fused):
    N, steps = (192, 6) if "Rough" in task else (384, 8)
    spec = make_spec(task, N)
    _, _, n_resets, _ = run_step_parity(spec, steps, fused=bool(fused), speculative=fused == "spec", tag=task)
    assert n_resets > 0, "test horizon produced no resets"


class _AsOracle:
    """FusedEnv under OracleEnv's attribute names, for tests.test_oracle_mdp_golden.compare_state."""
    _MAP = dict(cs_hist="cs_forces_hist", cs_last_air="cs_last_air_time", cs_cur_air="cs_current_air_time",
                cs_last_contact="cs_last_contact_time", cs_cur_contact="cs_current_contact_time", rc_hits="rc_ray_hits_w")

    def __init__(self, fus, spec):
        self.f, self.spec = fus, spec

    def __getattr__(self, name):
        f, spec = self.__dict__["f"], self.__dict__["spec"]
        if name == "term_dones":
            return f.term_dones[:len(spec["terminations"])].cpu()
        if name == "episode_sums":
            return f.episode_sums[:len(spec["rewards"])].cpu()
        if name == "step_reward":
            return f.step_reward[:, :len(spec["rewards"])].cpu()
        if name == "interval_time_left":
            return [t.cpu() for t in f.interval_time_left]
        if name == "interval_global_time":
            return {k: f.interval_timer(k).cpu() for k in f.interval_global_time}
        v = getattr(f, self._MAP.get(name, name))
        return v.cpu() if isinstance(v, torch.Tensor) else v


@pytest.mark.parametrize("fused", [True, False], ids=["fused-decimation", "per-substep"])
@pytest.mark.parametrize("task", TASKS)
def test_fused_step_matches_reference_golden(task, fused):
    """CUDA path against the fixtures recorded from the UNMODIFIED reference ManagerBasedRLEnv.step
    (tests/golden/gen_mdp.py) -- no oracle in between."""
    from isaaclab_b200.envs.fused_env import FusedEnv
    from isaaclab_b200.sim.synthetic import SyntheticSim
    from tests.test_oracle_mdp_golden import compare_state, load_fixture
    fx = load_fixture(task)
    N, steps = fx["N"], fx["steps"]
    spec = make_spec(task, N)
    fus = FusedEnv(spec, device="cuda")
    init_fused_state(fus, fx["init"])
    sim = SyntheticSim(spec, K=steps + 1, device="cuda", slabs={k: v.clone() for k, v in fx["slabs"].items()})
    view = _AsOracle(fus, spec)
    prev = sim.views(0)[-1]
    logs_seen, n_logs = {}, 0
    for k in range(steps):
        ref = fx["out"][k]
        ur, us_, uo = ref["u_reset"].cuda(), ref["u_step"].cuda(), ref["u_obs"].cuda()
        fus.set_uniforms(ur, us_, uo)
        obs, rew, term, trunc = fus.step(ref["action_in"].cuda(), sim.views(k + 1), prev, fused=fused)
        torch.cuda.synchronize()
        tag = f"{task} step {k} (vs reference)"
        compare_state(view, ref, spec, tag, rtol=1e-5, atol=2e-6, torque_tol=dict(rtol=2e-5, atol=2e-5), ray_tol=1e-4)
        close(sim.dev["root_state_w"][k + 1], ref["root_state_w"], tag + " root_state_w after events", atol=2e-6)
        close(sim.dev["joint_pos"][k + 1, -1], ref["joint_pos"], tag + " joint_pos after events")
        close(sim.dev["joint_vel"][k + 1, -1], ref["joint_vel"], tag + " joint_vel after events")
        close(obs, ref["obs"], tag + " obs", rtol=1e-5, atol=1e-4 if spec.get("raycaster") else 2e-6)
        if spec.get("raycaster"):
            D = fus.obs_dim - spec["R"]
            close(obs[:, :D], ref["obs"][:, :D], tag + " obs (proprioceptive part)", atol=2e-6)
        for st_f, st_r in zip(fus.act_state, ref["act_state"]):
            for key in st_r:
                close(st_f[key], st_r[key], tag + f" actuator state {key}", rtol=1e-4, atol=1e-5)
        # extras["log"] of this step (manager_based_rl_env.py:365-389 of the reference): same keys, same values
        log = fus.last_step_log()
        if int(ref["reset_buf"].sum()) == 0:   # the reference's extras["log"] then still holds an older step's dict
            assert log == {}
        else:
            assert set(log) == set(ref["log"]), f"{tag}: log keys {sorted(set(log) ^ set(ref['log']))} differ"
            for key, val in ref["log"].items():
                assert abs(log[key] - val) <= 1e-5 + 1e-5 * abs(val), f"{tag}: log[{key}] {log[key]} vs {val}"
                logs_seen[key] = logs_seen.get(key, 0.0) + val
            n_logs += 1
        prev = sim.views(k + 1)[-1]
    # the device-side mean over the steps that had resets == the mean of the reference's per-step dicts
    mean = fus.mean_log()
    assert n_logs > 0 and set(mean) == set(logs_seen)
    for key, tot in logs_seen.items():
        assert abs(mean[key] - tot / n_logs) <= 1e-5 + 1e-5 * abs(tot / n_logs), f"{task}: mean log[{key}]"
    assert fus.mean_log() == {}, "mean_log() did not reset the running sums"


@pytest.mark.parametrize("task", ["Isaac-Velocity-Rough-Anymal-C-v0", "Isaac-Cartpole-v0"])
def test_host_replay_prefetch_matches_device_replay(task):
    """ManagerBasedRLEnv.step over host-resident simulator state (pinned slabs, H2D prefetched one step ahead on a copy
    stream) against the same slabs resident on the device: identical bits, every step (the three staging sets rotate twice)."""
    from isaaclab_b200.envs.manager_based_rl_env import DeviceReplaySim, HostReplaySim, ManagerBasedRLEnv
    from isaaclab_b200.sim.synthetic import SyntheticSim, generate_slabs
    from isaaclab_b200.tasks import TASKS
    N, K = 256, 8  # no wrap: reset events write into the device-resident slabs, the host slabs stay pristine
    spec = make_spec(task, N)
    slabs = generate_slabs(spec, K, seed=77)
    g = torch.Generator().manual_seed(5)
    outs = []
    for kind in ("device", "host", "host-noprefetch", "host-pipelined", "host-graph"):
        if kind == "device":
            syn = SyntheticSim(spec, K=K, device="cuda", slabs={k: v.clone() for k, v in slabs.items()})
            sim = DeviceReplaySim(syn)
            sim.k = K - 1  # same "previous" slab as the host replay's first step

            def advance(sim=sim):  # cycle like the host replay does
                sim.k = (sim.k + 1) % K
                return sim.syn.views(sim.k)
            sim.advance = advance
        else:
            sim = HostReplaySim(spec, {k: v.clone() for k, v in slabs.items()}, device="cuda", prefetch=kind != "host-noprefetch")
        env = ManagerBasedRLEnv(TASKS[task](), sim=sim, device="cuda", spec=copy.deepcopy(spec), seed=3,
                                use_cuda_graph=kind == "host-graph")
        env.episode_length_buf[::5] = spec["max_episode_length"] - 3
        act_pinned = torch.empty(N, spec["A"]).pin_memory()
        g.manual_seed(5)
        rec = []
        slab = torch.empty(env.fused.result.numel(), dtype=torch.uint8).pin_memory()
        for _ in range(K - 1):
            act = torch.randn(N, spec["A"], generator=g)
            if kind in ("host-pipelined", "host-graph"):   # results land in a pinned slab; graph: captured per staging set
                act_pinned.copy_(act)
                env.step(act_pinned, results_out=slab)
                env.wait_results()
                v = env.result_views(slab)
                rec.append((v["obs"].clone().cuda(), v["reward"].clone().cuda(), v["terminated"].bool().cuda(),
                            v["truncated"].bool().cuda()))
                continue
            obs, rew, term, trunc, _ = env.step(act.pin_memory())
            rec.append((obs["policy"].clone(), rew.clone(), term.clone(), trunc.clone()))
        torch.cuda.synchronize()
        outs.append(rec)
    assert any(bool(r[3].any()) for r in outs[0]), "no time-outs in the horizon"
    for arm, other in zip(("host", "host-noprefetch", "host-pipelined", "host-graph"), outs[1:]):
        for k, (a, b) in enumerate(zip(outs[0], other)):
            for x, y, name in zip(a, b, ("obs", "reward", "terminated", "time_outs")):
                assert torch.equal(x, y), f"{task} step {k}: {name} differs between device and {arm} replay"


def test_observation_noise_kinds_match_oracle():
    """Every (constant | uniform | gaussian) x (add | scale | abs) noise model on the observation terms, CUDA against the
    oracle (itself pinned bitwise to the reference noise functions, tests/test_oracle_math.py) on the same draws."""
    from isaaclab_b200.envs.fused_env import FusedEnv
    from isaaclab_b200.sim.synthetic import SyntheticSim, generate_slabs
    task, N = "Isaac-Velocity-Flat-Anymal-C-v0", 128
    spec = make_spec(task, N)
    combos = [(k, a, b, op) for op in ("add", "scale", "abs")
              for k, a, b in (("constant", 0.37, 0.0), ("uniform", -0.1, 0.25), ("gaussian", 0.05, 0.3))]
    terms = spec["observations"]["terms"]
    for i, t in enumerate(terms):
        t["noise"] = combos[i % len(combos)]
    assert len(terms) >= 7
    slabs = generate_slabs(spec, 2, seed=3)
    orc = OracleEnv(copy.deepcopy(spec))
    fus = FusedEnv(spec, device="cuda")
    sim = SyntheticSim(spec, K=2, device="cuda", slabs={k: v.clone() for k, v in slabs.items()})
    g = torch.Generator().manual_seed(8)
    u_obs = torch.randn(N, fus.obs_dim, generator=g)   # gaussian columns read it as N(0,1), uniform columns as a draw
    fus.set_uniforms(None, None, u_obs.cuda())
    got = fus.compute_obs(sim.views(1)[-1]).cpu()
    want = orc.compute_obs(slab_views_cpu(slabs, 1)[-1], u_obs)
    close(got, want, "obs with mixed noise models", rtol=1e-5, atol=2e-6)
    # Philox path: gaussian columns must be standard-normal distributed around the clean value
    clean = copy.deepcopy(spec)
    for t in clean["observations"]["terms"]:
        t["noise"] = None
    for t in spec["observations"]["terms"]:
        t["noise"] = ("gaussian", 0.0, 1.0, "add")
    f_noisy, f_clean = FusedEnv(spec, device="cuda", seed=5), FusedEnv(clean, device="cuda", seed=5)
    z = (f_noisy.compute_obs(sim.views(1)[-1]) - f_clean.compute_obs(sim.views(1)[-1])).flatten()
    assert abs(float(z.mean())) < 0.05 and abs(float(z.std()) - 1.0) < 0.05, (float(z.mean()), float(z.std()))
    assert abs(float((z ** 3).mean())) < 0.15 and abs(float((z ** 4).mean()) - 3.0) < 0.3


@pytest.mark.parametrize("pattern", ["lidar", "bpearl"])
def test_height_scan_with_lidar_patterns_matches_oracle(pattern):
    """Non-vertical ray patterns (LAB/sensors/ray_caster/patterns/patterns.py:106-180) through the step: the observation
    kernel's general-ray instantiation (2-D DDA over the bin grid) against the oracle's brute-force query, incl. misses
    (inf -> clipped by the term) and a sensor that follows the full base orientation."""
    from isaaclab_b200.envs.compile import compile_env_cfg
    from isaaclab_b200.sensors import patterns as P
    from isaaclab_b200.tasks import TASKS as REG
    from tests.mdp_common import small_terrain
    cfg = REG["Isaac-Velocity-Rough-Anymal-C-v0"]()
    cfg.scene.terrain.max_init_terrain_level = 2
    hs = cfg.scene.height_scanner
    if pattern == "lidar":
        hs.pattern_cfg = P.LidarPatternCfg(channels=4, vertical_fov_range=(-60.0, -20.0), horizontal_fov_range=(-180.0, 180.0),
                                           horizontal_res=20.0)
        hs.attach_yaw_only = False
    else:
        hs.pattern_cfg = P.BpearlPatternCfg(horizontal_res=30.0, vertical_ray_angles=[-80.0, -60.0, -35.0, -10.0])   # negative: the rays of an upright mount point down
    hs.offset.pos = (0.0, 0.0, 2.0)     # the height scanner's 20 m mast would put most slanted rays off the terrain
    hs.max_distance = 8.0               # ... and the flattest ones still miss: both outcomes are covered
    spec = compile_env_cfg(cfg, num_envs=96, terrain=small_terrain(), seed=0)
    assert spec["R"] == (4 * 17 if pattern == "lidar" else 4 * 12)
    _, orc, n_resets, _ = run_step_parity(spec, 4, fused=True, tag=f"rough-anymal/{pattern}")
    hits = orc.rc_hits[..., 2]
    assert torch.isfinite(hits).any() and torch.isinf(hits).any() and n_resets > 0


def test_metrics_only_readback_matches_full_readback():
    """`ManagerBasedRLEnv.results_mode = "metrics"`: reward / terminated / time-outs / reset mask arrive in the host slab
    exactly as in the full read-back, the slab's observation part is not touched, the device observations are the same."""
    from isaaclab_b200.envs.manager_based_rl_env import HostReplaySim, ManagerBasedRLEnv
    from isaaclab_b200.sim.synthetic import generate_slabs
    from isaaclab_b200.tasks import TASKS as REG
    task, N = "Isaac-Velocity-Rough-Anymal-C-v0", 192
    spec = make_spec(task, N)
    slabs = generate_slabs(spec, 8, seed=3)
    outs = []
    for mode in ("all", "metrics"):
        env = ManagerBasedRLEnv(REG[task](), sim=HostReplaySim(spec, slabs, device="cuda"), device="cuda", spec=copy.deepcopy(spec),
                                seed=1, use_cuda_graph=True)
        env.results_mode = mode
        env.episode_length_buf[::5] = spec["max_episode_length"] - 3
        res = torch.full((env.fused.result.numel(),), 0x5A, dtype=torch.uint8).pin_memory()
        act = torch.randn(N, spec["A"], generator=torch.Generator().manual_seed(2)).pin_memory()
        rec = []
        for _ in range(6):
            env.step(act, results_out=res)
            env.wait_results()
            rec.append((res.clone(), env.fused.obs_buf.clone()))
        outs.append(rec)
    o_rew = env.fused.result_layout["reward"][0]
    for k, ((full, obs_f), (met, obs_m)) in enumerate(zip(*outs)):
        assert torch.equal(full[o_rew:], met[o_rew:]), f"step {k}: metrics differ"
        assert bool((met[:o_rew] == 0x5A).all()), f"step {k}: observation part of the slab was written"
        assert torch.equal(obs_f, obs_m), f"step {k}: device observations differ"
