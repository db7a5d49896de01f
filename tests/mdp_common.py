"""Shared helpers of the MDP parity tests: small task specs, seeded uniforms, state comparison."""
import copy

import numpy as np
import torch

from isaaclab_b200 import terrains
from isaaclab_b200.envs.compile import compile_env_cfg
from isaaclab_b200.sim.synthetic import generate_slabs, initial_env_state
from isaaclab_b200.tasks import TASKS

_TERRAIN = {}


def small_terrain():
    if "t" not in _TERRAIN:
        cfg = copy.deepcopy(terrains.ROUGH_TERRAINS_CFG)
        cfg.num_rows, cfg.num_cols, cfg.border_width = 3, 6, 4.0
        _TERRAIN["t"] = terrains.TerrainGenerator(cfg)
    return _TERRAIN["t"]


def make_spec(task, N, seed=0):
    cfg = TASKS[task]()
    terr = None
    if cfg.scene.terrain is not None and cfg.scene.terrain.terrain_type == "generator":
        terr = small_terrain()
        cfg.scene.terrain.max_init_terrain_level = 2
    return compile_env_cfg(cfg, num_envs=N, terrain=terr, seed=seed)


def draw_uniforms(n_steps, N, n_reset, n_step, n_obs, seed):
    g = torch.Generator().manual_seed(seed)
    return [(torch.rand(N, n_reset, generator=g), torch.rand(N, n_step, generator=g), torch.rand(N, n_obs, generator=g))
            for _ in range(n_steps)]


def slab_views_cpu(slabs, k):
    S = slabs["joint_pos"].shape[1]
    return [dict(root_state_w=slabs["root_state_w"][k], joint_pos=slabs["joint_pos"][k, s], joint_vel=slabs["joint_vel"][k, s],
                 net_forces_w=slabs["net_forces_w"][k, s], body_lin_vel_w=slabs["body_lin_vel_w"][k]) for s in range(S)]


def init_oracle_state(env, init):
    env.episode_length_buf[:] = init["episode_length_buf"]
    if env.spec.get("command"):
        env.cmd_time_left[:] = init["cmd_time_left"]
    for t in env.interval_time_left:
        t[:] = init["interval_time_left"]
    for gt in env.interval_global_time.values():
        gt[:] = init.get("interval_global_time_left", 0.03)
    from isaaclab_b200.sim.synthetic import initial_delay_lags
    for act, st in zip(env.spec["actuators"], env.act_state):
        if act["kind"] in ("delayed_pd", "remotized_pd"):
            st["lags"][:] = initial_delay_lags(act, init)


def init_fused_state(env, init):
    from isaaclab_b200.sim.synthetic import init_fused_state as _init
    _init(env, init)


_FULL_TERRAIN = {}


def full_terrain():
    """The 10 x 20 `ROUGH_TERRAINS_CFG` mesh the bench times (1.19 M triangles)."""
    if "t" not in _FULL_TERRAIN:
        _FULL_TERRAIN["t"] = terrains.TerrainGenerator(copy.deepcopy(terrains.ROUGH_TERRAINS_CFG))
    return _FULL_TERRAIN["t"]


def make_full_spec(task, N, seed=0):
    """Spec of `task` at its BASELINE size on the full terrain, exactly as bench.py builds it."""
    cfg = TASKS[task]()
    terr = None
    if cfg.scene.terrain is not None and cfg.scene.terrain.terrain_type == "generator":
        terr = full_terrain()
    return compile_env_cfg(cfg, num_envs=N, terrain=terr, seed=seed)


def close(a, b, name, rtol=1e-5, atol=1e-6):
    torch.testing.assert_close(a.detach().cpu().float(), b.float(), rtol=rtol, atol=atol, msg=lambda m: f"{name}: {m}")


def exact(a, b, name):
    a = a.detach().cpu()
    assert torch.equal(a.to(b.dtype), b), f"{name}: {(a.to(b.dtype) != b).sum().item()} mismatches"


def compare_fused_vs_oracle(fus, orc, spec, tag, f_out, o_out):
    """Every buffer the step leaves behind, CUDA path vs oracle: masks / counters / resample indices / terrain levels
    bit-exact; floats rtol 1e-5, atol 1e-6..2e-6; torques 2e-5 (SFU gate activations); ray hits 1e-4 m."""
    f_obs, f_rew, f_term, f_trunc = f_out
    o_obs, o_rew, o_term, o_trunc = o_out
    exact(f_term, o_term, tag + " terminated")
    exact(f_trunc, o_trunc, tag + " truncated")
    exact(fus.reset_buf, orc.reset_buf, tag + " reset_buf")
    exact(fus.episode_length_buf, orc.episode_length_buf, tag + " episode_length_buf")
    exact(fus.term_dones[:len(spec["terminations"])], orc.term_dones, tag + " term_dones")
    close(f_rew, o_rew, tag + " reward")
    close(fus.step_reward[:, :len(spec["rewards"])], orc.step_reward, tag + " step_reward", atol=2e-6)
    close(fus.episode_sums[:len(spec["rewards"])], orc.episode_sums, tag + " episode_sums")
    close(fus.computed_torque, orc.computed_torque, tag + " computed_torque", rtol=2e-5, atol=2e-5)
    close(fus.applied_torque, orc.applied_torque, tag + " applied_torque", rtol=2e-5, atol=2e-5)
    close(fus.joint_acc, orc.joint_acc, tag + " joint_acc", atol=1e-4)
    close(fus.action, orc.action, tag + " action")
    close(fus.prev_action, orc.prev_action, tag + " prev_action")
    if spec.get("command"):
        exact(fus.cmd_counter, orc.cmd_counter, tag + " command_counter (resample indices)")
        exact(fus.cmd_is_heading, orc.cmd_is_heading, tag + " is_heading_env")
        exact(fus.cmd_is_standing, orc.cmd_is_standing, tag + " is_standing_env")
        close(fus.cmd_time_left, orc.cmd_time_left, tag + " time_left")
        close(fus.cmd_vel_b, orc.cmd_vel_b, tag + " vel_command_b", atol=2e-6)
        close(fus.cmd_heading_target, orc.cmd_heading_target, tag + " heading_target")
        close(fus.cmd_metric_xy, orc.cmd_metric_xy, tag + " metric xy")
    if spec.get("contact"):
        close(fus.cs_forces_hist, orc.cs_hist, tag + " force history")
        for a, b in (("cs_last_air_time", "cs_last_air"), ("cs_current_air_time", "cs_cur_air"),
                     ("cs_last_contact_time", "cs_last_contact"), ("cs_current_contact_time", "cs_cur_contact")):
            close(getattr(fus, a), getattr(orc, b), tag + " " + a)
    if spec.get("curriculum"):
        exact(fus.terrain_levels, orc.terrain_levels, tag + " terrain_levels")
        close(fus.env_origins, orc.env_origins, tag + " env_origins")
    for i, t in enumerate(fus.interval_time_left):
        if i in orc.interval_global_time:   # is_global_time: one timer (the fused path double-buffers it by step parity)
            close(fus.interval_timer(i), orc.interval_global_time[i], tag + " global interval time_left")
        else:
            close(t, orc.interval_time_left[i], tag + " interval time_left")
    if spec.get("raycaster"):
        if getattr(fus, "rc_ray_hits_w", None) is not None:
            hz_f, hz_o = fus.rc_ray_hits_w[..., 2].cpu(), orc.rc_hits[..., 2]
            bad = ~(((hz_f - hz_o).abs() <= 1e-4) | (torch.isinf(hz_f) & torch.isinf(hz_o)))
            assert bad.sum().item() == 0, f"{tag}: {bad.sum().item()} ray hits differ by more than 1e-4 m"
        close(fus.rc_drift, orc.rc_drift, tag + " drift")
    close(f_obs, o_obs, tag + " obs", rtol=1e-5, atol=1e-4 if spec.get("raycaster") else 2e-6)
    if spec.get("raycaster"):
        D = fus.obs_dim - spec["R"]
        close(f_obs[:, :D], o_obs[:, :D], tag + " obs (proprioceptive part)", atol=2e-6)


def run_step_parity(spec, steps, fused=True, speculative=False, oracle_grid=None, force_timeouts=7, tag="",
                    seeds=(1234, 5, 99, 3), fused_kwargs=None):
    """CUDA path (through the C ABI) against the CPU oracle on the same seeded simulator state and pre-drawn uniforms for
    `steps` env steps; returns (fused env, oracle env, resets, command resamples).  `oracle_grid`: cell size of the
    oracle's exact xy pre-filter for vertical rays (needed at full terrain size; identical results to brute force,
    tests/test_oracle_raycast.py)."""
    from isaaclab_b200.envs.fused_env import FusedEnv
    from isaaclab_b200.sim.synthetic import SyntheticSim
    from oracle.mdp_env import OracleEnv
    N = spec["N"]
    slabs = generate_slabs(spec, steps + 1, seed=seeds[0])
    init = initial_env_state(spec, seed=seeds[1])
    if force_timeouts:
        init["episode_length_buf"][::force_timeouts] = spec["max_episode_length"] - 2  # time-outs within the horizon
    orc = OracleEnv(copy.deepcopy(spec))
    if oracle_grid and spec.get("raycaster"):
        orc.mesh.enable_grid(oracle_grid)
    fus = FusedEnv(spec, device="cuda", **(fused_kwargs or {}))
    fus.speculative_scan = bool(speculative)
    init_oracle_state(orc, init)
    init_fused_state(fus, init)
    sim = SyntheticSim(spec, K=steps + 1, device="cuda", slabs={k: v.clone() for k, v in slabs.items()})
    us = draw_uniforms(steps, N, fus.n_u_reset, fus.n_u_step, fus.obs_dim, seed=seeds[2])
    g = torch.Generator().manual_seed(seeds[3])
    n_resets = n_resample = 0
    prev_cpu, prev_gpu = slab_views_cpu(slabs, 0)[-1], sim.views(0)[-1]
    for k in range(steps):
        action = torch.randn(N, spec["A"], generator=g)
        u_r, u_s, u_o = us[k]
        views_cpu = slab_views_cpu(slabs, k + 1)
        o_out = orc.step(action, views_cpu, u_r, u_s, u_o, prev_cpu)
        ur, us_, uo = u_r.cuda(), u_s.cuda(), u_o.cuda()
        fus.set_uniforms(ur, us_, uo)
        f_out = fus.step(action.cuda(), sim.views(k + 1), prev_gpu, fused=fused)
        torch.cuda.synchronize()
        t = f"{tag} step {k}"
        compare_fused_vs_oracle(fus, orc, spec, t, f_out, o_out)
        # simulator state written by reset / push events
        close(sim.dev["root_state_w"][k + 1], views_cpu[-1]["root_state_w"], t + " root_state_w after events", atol=2e-6)
        close(sim.dev["joint_pos"][k + 1, -1], views_cpu[-1]["joint_pos"], t + " joint_pos after events")
        close(sim.dev["joint_vel"][k + 1, -1], views_cpu[-1]["joint_vel"], t + " joint_vel after events")
        if spec.get("command"):
            n_resample += len(orc.last_resample_ids)
        n_resets += int(orc.reset_buf.sum())
        prev_cpu, prev_gpu = views_cpu[-1], sim.views(k + 1)[-1]
    for st_f, st_o in zip(fus.act_state, orc.act_state):
        if "ring" in st_f:   # DelayedPD: same lags, and the same set-points at every lag (the ring layouts differ:
            # the CUDA path keeps one write pointer per joint, the reference one per buffer)
            assert torch.equal(st_f["lags"].cpu().long(), st_o["lags"]), f"{tag} delay lags"
            ring_f, cnt = st_f["ring"].cpu(), st_f["count"].cpu().long()          # [N, Jg, L, 3], [N, Jg]
            L = ring_f.shape[2]
            for lag in range(L):
                k = torch.minimum(torch.full_like(cnt, lag), cnt - 1).clamp(min=0)
                got = torch.gather(ring_f, 2, ((cnt - 1 - k) % L)[..., None, None].expand(-1, -1, 1, 3)).squeeze(2)
                want = st_o["ring"].lagged(torch.full((fus.N,), lag))
                close(got, want, f"{tag} delayed set-points at lag {lag}", rtol=1e-6, atol=1e-7)
            continue
        for key in st_f:
            close(st_f[key], st_o[key], f"{tag} actuator state {key}", rtol=1e-4, atol=1e-5)
    return fus, orc, n_resets, n_resample
