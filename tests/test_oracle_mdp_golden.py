"""Pins oracle/mdp_env.py against fixtures produced by the UNMODIFIED reference `ManagerBasedRLEnv.step`
(tests/golden/gen_mdp.py; real managers, real task cfgs, real actuator / sensor / articulation-data classes over replayed
PhysX views).  The oracle replays the recorded uniforms and must reproduce every output and internal buffer.

Both sides are torch fp32 on the CPU, so the float bar here is tighter than the GPU one; masks, counters and indices are
exact."""
import copy
import os

import pytest
import torch

from oracle.mdp_env import OracleEnv
from tests.mdp_common import init_oracle_state, make_spec, slab_views_cpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TASKS = ["Isaac-Cartpole-v0", "Isaac-Velocity-Flat-Anymal-C-v0", "Isaac-Velocity-Rough-Anymal-C-v0",
         "Isaac-Velocity-Rough-Unitree-Go1-v0", "Isaac-Velocity-Rough-H1-v0",
         "B200-Opcode-Coverage-v0"]   # synthetic cfg: the opcodes / actuator kinds / options the five tasks do not use


def fixture_path(task):
    return os.path.join(GOLDEN, "mdp_" + task.replace("Isaac-", "").replace("-", "_").lower() + ".pt")


def load_fixture(task):
    return torch.load(fixture_path(task), weights_only=False)


def close(a, b, name, rtol=1e-6, atol=1e-6):
    torch.testing.assert_close(a.float(), b.float(), rtol=rtol, atol=atol, msg=lambda m: f"{name}: {m}")


def exact(a, b, name):
    assert torch.equal(a.to(b.dtype), b), f"{name}: {(a.to(b.dtype) != b).sum().item()} mismatches"


def compare_state(env, ref, spec, tag, rtol=1e-6, atol=1e-6, torque_tol=None, ray_tol=1e-6):
    """`env`: an object with OracleEnv's attribute names (the CUDA env is adapted to it in the GPU test)."""
    tt = torque_tol or dict(rtol=rtol, atol=1e-5)
    exact(env.reset_buf, ref["reset_buf"], tag + " reset_buf")
    exact(env.terminated, ref["terminated"], tag + " terminated")
    exact(env.truncated, ref["time_outs"], tag + " time_outs")
    exact(env.episode_length_buf, ref["episode_length_buf"], tag + " episode_length_buf")
    exact(env.term_dones, ref["term_dones"], tag + " term_dones")
    close(env.reward_buf, ref["reward"], tag + " reward", rtol, atol)
    close(env.step_reward, ref["step_reward"], tag + " step_reward", rtol, 2 * atol)
    close(env.episode_sums, ref["episode_sums"], tag + " episode_sums", rtol, atol)
    close(env.action, ref["action"], tag + " action", rtol, atol)
    close(env.prev_action, ref["prev_action"], tag + " prev_action", rtol, atol)
    close(env.joint_pos_target, ref["joint_pos_target"], tag + " joint_pos_target", rtol, atol)
    close(env.computed_torque, ref["computed_torque"], tag + " computed_torque", **tt)
    close(env.applied_torque, ref["applied_torque"], tag + " applied_torque", **tt)
    close(env.joint_acc, ref["joint_acc"], tag + " joint_acc", rtol, 1e-4)
    if spec.get("command"):
        exact(env.cmd_counter, ref["cmd_counter"], tag + " command_counter")
        exact(env.cmd_is_heading, ref["cmd_is_heading"], tag + " is_heading_env")
        exact(env.cmd_is_standing, ref["cmd_is_standing"], tag + " is_standing_env")
        close(env.cmd_time_left, ref["cmd_time_left"], tag + " time_left", rtol, atol)
        close(env.cmd_vel_b, ref["cmd_vel_b"], tag + " vel_command_b", rtol, 2 * atol)
        close(env.cmd_heading_target, ref["cmd_heading_target"], tag + " heading_target", rtol, atol)
        close(env.cmd_metric_xy, ref["cmd_metric_xy"], tag + " metric xy", rtol, atol)
        close(env.cmd_metric_yaw, ref["cmd_metric_yaw"], tag + " metric yaw", rtol, atol)
    if spec.get("contact"):
        close(env.cs_hist, ref["cs_hist"], tag + " net_forces_w_history", rtol, atol)
        for k in ("cs_last_air", "cs_cur_air", "cs_last_contact", "cs_cur_contact"):
            close(getattr(env, k), ref[k], tag + " " + k, rtol, atol)
    if spec.get("curriculum"):
        exact(env.terrain_levels, ref["terrain_levels"], tag + " terrain_levels")
    close(env.env_origins, ref["env_origins"], tag + " env_origins", rtol, atol)
    for i, t in enumerate(env.interval_time_left):
        if i in env.interval_global_time:   # is_global_time: the reference keeps a [1] tensor (event_manager.py:368-372)
            t = env.interval_global_time[i]
        close(t, ref["interval_time_left"][i], tag + f" interval[{i}] time_left", rtol, atol)
    if spec.get("raycaster"):
        close(env.rc_drift, ref["rc_drift"], tag + " drift", rtol, atol)
        hz_a, hz_b = env.rc_hits[..., 2].float().cpu(), ref["rc_hits"][..., 2]
        bad = ~(((hz_a - hz_b).abs() <= ray_tol) | (torch.isinf(hz_a) & torch.isinf(hz_b)))
        assert bad.sum().item() == 0, f"{tag}: {bad.sum().item()} ray hit heights differ by more than {ray_tol}"


@pytest.mark.parametrize("task", TASKS)
def test_oracle_reproduces_reference_step(task):
    if not os.path.exists(fixture_path(task)):
        pytest.fail(f"golden fixture missing: {fixture_path(task)} (python -m tests.golden.gen_mdp)")
    fx = load_fixture(task)
    N, steps = fx["N"], fx["steps"]
    spec = make_spec(task, N)
    slabs = {k: v.clone() for k, v in fx["slabs"].items()}
    orc = OracleEnv(copy.deepcopy(spec))
    init_oracle_state(orc, fx["init"])
    prev = slab_views_cpu(slabs, 0)[-1]
    n_resets = 0
    for k in range(steps):
        ref = fx["out"][k]
        views = slab_views_cpu(slabs, k + 1)
        obs, rew, term, trunc = orc.step(ref["action_in"], views, ref["u_reset"], ref["u_step"], ref["u_obs"], prev)
        tag = f"{task} step {k}"
        compare_state(orc, ref, spec, tag)
        close(views[-1]["root_state_w"], ref["root_state_w"], tag + " root_state_w after events", 1e-6, 2e-6)
        close(views[-1]["joint_pos"], ref["joint_pos"], tag + " joint_pos after events")
        close(views[-1]["joint_vel"], ref["joint_vel"], tag + " joint_vel after events")
        close(obs, ref["obs"], tag + " obs", 1e-6, 2e-6)
        for st_o, st_r in zip(orc.act_state, ref["act_state"]):
            for key in st_r:
                close(st_o[key], st_r[key], tag + f" actuator state {key}", 1e-5, 1e-6)
        for key, val in ref["log"].items():
            assert key in orc.log, f"{tag}: oracle log lacks '{key}'"
            assert abs(float(orc.log[key]) - val) <= 1e-5 + 1e-5 * abs(val), f"{tag}: log[{key}] {float(orc.log[key])} vs {val}"
        n_resets += int(ref["reset_buf"].sum())
        prev = views[-1]
    assert n_resets > 0
