"""`FusedEnv`: the device-side state of one ManagerBasedRLEnv instance and its five fused kernels.

Input is the neutral env `spec` (plain dict; DESIGN.md "Env spec") that `isaaclab_b200.envs.compile` produces from
the reference-style manager cfg objects.  This module owns every tensor the reference keeps in ArticulationData,
ContactSensorData, RayCasterData and the managers (same shapes/dtypes, env-major), fills the `B200LabEnvState`
descriptor (include/b200lab.h) with their raw pointers once, and launches

    process_action -> S x (substep_pre, [simulator], substep_post) -> post_step -> post_reset_write -> compute_obs

through the C ABI on torch's current stream.  Nothing here computes on the host or falls back to torch ops.
"""
from __future__ import annotations

import ctypes
import math

import numpy as np
import torch

from .. import _lib

C = _lib.consts
f32 = np.float32


def _f32span_tensor(lo, hi):
    """`upper - lower` as the reference computes it on fp32 tensors / inside Tensor.uniform_."""
    return float(f32(hi) - f32(lo))


def _f32span_py(lo, hi):
    """`upper - lower` computed in python doubles, then used as an fp32 scalar (math.py:1331, noise_model.py:64)."""
    return float(f32(float(hi) - float(lo)))


_ACT_KIND = {"implicit": "B200LAB_ACT_IMPLICIT", "ideal_pd": "B200LAB_ACT_IDEAL_PD", "dc_motor": "B200LAB_ACT_DC_MOTOR",
             "delayed_pd": "B200LAB_ACT_DELAYED_PD", "remotized_pd": "B200LAB_ACT_REMOTIZED_PD",
             "lstm": "B200LAB_ACT_NET_LSTM", "mlp": "B200LAB_ACT_NET_MLP"}
_MLP_ACT = {"softsign": 0, "elu": 1, "relu": 2, "tanh": 3}
_KEYS6 = ("x", "y", "z", "roll", "pitch", "yaw")


def _mask(ids):
    m = 0
    for i in ids:
        m |= 1 << int(i)
    return m


class SimView:
    """Raw views of the simulator's state tensors for one physics step (B200LabSimView)."""

    def __init__(self, root_state_w, joint_pos, joint_vel, net_forces_w=None, body_lin_vel_w=None):
        self.root_state_w, self.joint_pos, self.joint_vel = root_state_w, joint_pos, joint_vel
        self.net_forces_w, self.body_lin_vel_w = net_forces_w, body_lin_vel_w
        v = _lib.structs["B200LabSimView"]()
        v.root_state_w, v.joint_pos, v.joint_vel = _lib.ptr(root_state_w), _lib.ptr(joint_pos), _lib.ptr(joint_vel)
        v.net_forces_w, v.body_lin_vel_w = _lib.ptr(net_forces_w), _lib.ptr(body_lin_vel_w)
        self.c = v
        self.ref = ctypes.byref(v)


class FusedEnv:
    def __init__(self, spec: dict, device="cuda", seed: int = 0, store_ray_hits: bool = True):
        """`store_ray_hits=False`: RayCasterData.ray_hits_w [N,R,3] is not materialised -- the hit heights go straight from
        registers into the observation row.  Only valid for a sensor that is outdated at every observation
        (update_period <= env step dt), which is checked here."""
        self.store_ray_hits = bool(store_ray_hits)
        self.spec = s = spec
        self.device = dev = torch.device(device)
        self.L = _lib.lib()
        N, J, A, B, R = s["N"], s["J"], s["A"], s["B"], s.get("R", 0)
        if J > 64 or B > 64:
            raise ValueError("joint/body selections are 64-bit masks: at most 64 joints and 64 bodies")
        self.N, self.J, self.A, self.B, self.R = N, J, A, B, R
        self._keep = []  # tensors referenced only through raw pointers
        z = lambda *shape, dtype=torch.float32: torch.zeros(*shape, dtype=dtype, device=dev)  # noqa: E731

        def const(x, dtype=torch.float32):
            t = torch.as_tensor(np.asarray(x), dtype=dtype).to(dev).contiguous()
            self._keep.append(t)
            return t

        st = _lib.structs["B200LabEnvState"]()
        p = _lib.ptr
        st.N, st.J, st.A, st.B, st.R = N, J, A, B, R
        # --- articulation
        self.default_root_state = const(s["default_root_state"])
        self.default_joint_pos = const(s["default_joint_pos"])
        self.default_joint_vel = const(s["default_joint_vel"])
        self.soft_joint_pos_limits = const(s["soft_joint_pos_limits"])
        self.soft_joint_vel_limits = const(s["soft_joint_vel_limits"])
        st.default_root_state, st.default_joint_pos = p(self.default_root_state), p(self.default_joint_pos)
        st.default_joint_vel, st.soft_joint_pos_limits = p(self.default_joint_vel), p(self.soft_joint_pos_limits)
        st.soft_joint_vel_limits = p(self.soft_joint_vel_limits)
        for name in ("joint_pos_target", "joint_vel_target", "joint_effort_target", "computed_torque", "applied_torque",
                     "joint_pos_target_sim", "joint_effort_target_sim", "prev_joint_vel", "joint_acc"):
            t = z(N, J)
            setattr(self, name, t)
            setattr(st, name, p(t))
        # --- action manager
        for name in ("action", "prev_action", "raw_actions", "processed_actions"):
            t = z(N, A)
            setattr(self, name, t)
            setattr(st, name, p(t))
        a = s["action"]
        scale = np.broadcast_to(np.asarray(a["scale"], dtype=f32), (A,))
        offset = np.broadcast_to(np.asarray(a["offset"], dtype=f32), (A,))
        st.action_scale, st.action_offset = p(const(scale.copy())), p(const(offset.copy()))
        st.action_clip = p(const(a["clip"])) if a.get("clip") is not None else None
        st.action_joint_ids = p(const(list(a["joint_ids"]), torch.int32))
        st.action_kind = 0 if a["kind"] == "joint_position" else 1
        for j in range(64):
            st.joint_action_col[j] = -1
            st.joint_group[j] = -1
            st.joint_local[j] = -1
        for col, j in enumerate(a["joint_ids"]):
            st.joint_action_col[int(j)] = col
        # --- actuators
        acts = s["actuators"]
        if len(acts) > C["B200LAB_MAX_ACT_GROUPS"]:
            raise ValueError("too many actuator groups")
        st.num_act_groups = len(acts)
        self.act_state = []
        for gi, act in enumerate(acts):
            g = st.act[gi]
            ids = [int(j) for j in act["joint_ids"]]
            Jg = len(ids)
            g.kind = C[_ACT_KIND[act["kind"]]]
            g.joint_mask = _mask(ids)
            g.group_joints = Jg
            for local, j in enumerate(ids):
                st.joint_group[j] = gi
                st.joint_local[j] = local

            def per_joint(val, default=0.0):
                full = np.zeros(J, dtype=f32)
                v = np.broadcast_to(np.asarray(default if val is None else val, dtype=f32), (Jg,))
                full[ids] = v
                return p(const(full))

            g.stiffness, g.damping = per_joint(act.get("stiffness")), per_joint(act.get("damping"))
            g.effort_limit = per_joint(act.get("effort_limit"), np.inf)
            g.velocity_limit = per_joint(act.get("velocity_limit"), np.inf)
            sat = act.get("saturation_effort")
            g.saturation_effort = float("inf") if sat is None else float(sat)
            stt = {}
            if act["kind"] == "lstm":
                net = act["lstm"]
                Lh, H = net["num_layers"], net["hidden"]
                parts = []
                for l in range(Lh):
                    parts += [net[f"weight_ih_l{l}"], net[f"weight_hh_l{l}"], net[f"bias_ih_l{l}"], net[f"bias_hh_l{l}"]]
                parts += [net["head_weight"], net["head_bias"]]
                w = torch.cat([torch.as_tensor(x, dtype=torch.float32).reshape(-1) for x in parts])
                g.net_weights = p(const(w.numpy()))
                g.lstm_layers, g.lstm_hidden = Lh, H
                stt["h"], stt["c"] = z(Lh, N * Jg, H), z(Lh, N * Jg, H)
                g.lstm_h, g.lstm_c = p(stt["h"]), p(stt["c"])
            elif act["kind"] == "mlp":
                net = act["mlp"]
                dims = [net["layers"][0][0].shape[1]] + [W.shape[0] for W, _ in net["layers"]]
                if len(dims) > 6:
                    raise ValueError("ActuatorNetMLP: at most 5 layers")
                parts = []
                for W, b in net["layers"]:
                    parts += [W, b]
                w = torch.cat([torch.as_tensor(x, dtype=torch.float32).reshape(-1) for x in parts])
                g.net_weights = p(const(w.numpy()))
                g.mlp_layers = len(net["layers"])
                for i, dd in enumerate(dims):
                    g.mlp_dims[i] = int(dd)
                g.mlp_act = _MLP_ACT[net["activation"]]
                hist = max(net["input_idx"]) + 1
                g.mlp_hist, g.mlp_num_idx = hist, len(net["input_idx"])
                for i, ii in enumerate(net["input_idx"]):
                    g.mlp_input_idx[i] = int(ii)
                g.mlp_vel_pos_order = 0 if net["input_order"] == "pos_vel" else 1
                g.mlp_pos_scale, g.mlp_vel_scale = float(net["pos_scale"]), float(net["vel_scale"])
                g.mlp_torque_scale = float(net["torque_scale"])
                stt["pos_hist"], stt["vel_hist"] = z(N, hist, Jg), z(N, hist, Jg)
                g.mlp_pos_hist, g.mlp_vel_hist = p(stt["pos_hist"]), p(stt["vel_hist"])
            elif act["kind"] in ("delayed_pd", "remotized_pd"):
                g.min_delay, g.max_delay = int(act["min_delay"]), int(act["max_delay"])
                k = sum(1 for a_ in acts[:gi] if a_["kind"] in ("delayed_pd", "remotized_pd"))
                g.delay_u_slot = C["B200LAB_U_RESET_JOINT"] + 2 * J * max(1, len(s.get("events", {}).get("reset_joints", []))) + k
                stt["ring"] = z(N, Jg, g.max_delay + 1, 3)
                stt["count"], stt["lags"] = z(N, Jg, dtype=torch.int32), z(N, dtype=torch.int32)
                g.delay_ring, g.delay_count, g.delay_lag = p(stt["ring"]), p(stt["count"]), p(stt["lags"])
                if act["kind"] == "remotized_pd":
                    g.lookup_n = len(act["lookup_angle"])
                    g.lookup_x, g.lookup_y = p(const(act["lookup_angle"])), p(const(act["lookup_torque"]))
            self.act_state.append(stt)
        self._st = st
        self._finish_init(seed)

    # ------------------------------------------------------------------------------------------------
    def _term(self, dst, kind, t):
        """Fill one B200LabTerm from a spec term dict."""
        prefix = {"rew": "B200LAB_REW_", "term": "B200LAB_TERM_", "obs": "B200LAB_OBS_"}[kind]
        if t["op"] == "external":   # user term evaluated in Python: its value enters through ext_values[row .. row+dim)
            dst.op = C[prefix + "EXTERNAL"]
            dst.weight = float(t.get("weight", 0.0))
            dst.time_out = 1 if t.get("time_out", False) else 0
            dst.i0 = int(t["ext_row"])
            return
        key = prefix + t["op"].upper()
        if key not in C:
            raise ValueError(f"term '{t['op']}' has no fused opcode ({key})")
        dst.op = C[key]
        dst.weight = float(t.get("weight", 0.0))
        dst.time_out = 1 if t.get("time_out", False) else 0
        dst.joint_mask = _mask(t["joint_ids"]) if t.get("joint_ids") is not None else _mask(range(self.J))
        dst.body_mask = _mask(t["body_ids"]) if t.get("body_ids") is not None else 0
        dst.body_mask2 = _mask(t["asset_body_ids"]) if t.get("asset_body_ids") is not None else 0
        op = t["op"]
        if "std" in t:
            dst.p[0] = float(t["std"]) ** 2
        if "threshold" in t:
            dst.p[0] = float(t["threshold"])
        if op == "feet_air_time":
            dst.p[1] = float(self.spec["step_dt"]) + 1.0e-8
        if op == "joint_pos_target_l2":
            dst.p[0] = float(t["target"])
        if op == "joint_pos_out_of_manual_limit":
            dst.p[0], dst.p[1] = float(t["bounds"][0]), float(t["bounds"][1])
        if op == "root_height_below_minimum":
            dst.p[0] = float(t["minimum_height"])
        if op == "bad_orientation":
            dst.p[0] = float(t["limit_angle"])
        if op == "height_scan":
            dst.p[0] = float(t.get("offset", 0.5))
        if op in ("feet_slide",) and bin(dst.body_mask).count("1") != bin(dst.body_mask2).count("1"):
            raise ValueError("feet_slide: sensor and asset body selections must have the same length")

    def _finish_init(self, seed):
        s, st, dev, p = self.spec, self._st, self.device, _lib.ptr
        N, J, A, B, R = self.N, self.J, self.A, self.B, self.R
        z = lambda *shape, dtype=torch.float32: torch.zeros(*shape, dtype=dtype, device=dev)  # noqa: E731

        def const(x, dtype=torch.float32):
            t = torch.as_tensor(np.asarray(x), dtype=dtype).to(dev).contiguous()
            self._keep.append(t)
            return t

        # --- contact sensor
        cs = s.get("contact")
        st.hist = 1
        if cs:
            H = int(cs["history_length"])
            if H < 1:
                raise ValueError("contact sensor: the fused path keeps a force history (history_length >= 1)")
            st.hist = H
            self.cs_net_forces_w, self.cs_forces_hist = z(N, B, 3), z(N, H, B, 3)
            st.cs_net_forces_w, st.cs_forces_hist = p(self.cs_net_forces_w), p(self.cs_forces_hist)
            if cs.get("track_air_time", True):
                for name in ("cs_last_air_time", "cs_current_air_time", "cs_last_contact_time", "cs_current_contact_time"):
                    t = z(N, B)
                    setattr(self, name, t)
                    setattr(st, name, p(t))
            self.cs_timestamp, self.cs_timestamp_last_update = z(N), z(N)
            self.cs_is_outdated = torch.ones(N, dtype=torch.uint8, device=dev)
            st.cs_timestamp, st.cs_timestamp_last_update = p(self.cs_timestamp), p(self.cs_timestamp_last_update)
            st.cs_is_outdated = p(self.cs_is_outdated)
            st.cs_force_threshold, st.cs_update_period = float(cs["force_threshold"]), float(cs["update_period"])
        # --- ray caster
        r = s.get("raycaster")
        self.mesh = None
        if r:
            self.rc_timestamp, self.rc_timestamp_last_update = z(N), z(N)
            self.rc_is_outdated = torch.ones(N, dtype=torch.uint8, device=dev)
            if not self.store_ray_hits and float(r["update_period"]) > float(s["step_dt"]) + 1e-6:
                raise ValueError("store_ray_hits=False needs a ray caster that updates at every observation "
                                 "(update_period <= env step dt): a sensor that skips updates re-reads its stored hits")
            self.rc_drift, self.rc_pos_w, self.rc_quat_w = z(N, 3), z(N, 3), z(N, 4)
            self.rc_ray_hits_w = z(N, R, 3) if self.store_ray_hits else None
            st.rc_timestamp, st.rc_timestamp_last_update = p(self.rc_timestamp), p(self.rc_timestamp_last_update)
            st.rc_is_outdated, st.rc_drift = p(self.rc_is_outdated), p(self.rc_drift)
            st.rc_pos_w, st.rc_quat_w, st.rc_ray_hits_w = p(self.rc_pos_w), p(self.rc_quat_w), p(self.rc_ray_hits_w)
            st.rc_ray_starts, st.rc_ray_dirs = p(const(r["ray_starts"])), p(const(r["ray_dirs"]))
            st.rc_update_period, st.rc_max_distance = float(r["update_period"]), float(r["max_distance"])
            st.rc_attach_yaw_only = 1 if r["attach_yaw_only"] else 0
            lo, hi = r["drift_range"]
            st.rc_drift_lo, st.rc_drift_span = float(lo), _f32span_tensor(lo, hi)
            from ..sensors.ray_caster import WarpFreeMesh
            self.mesh = r["mesh"] if isinstance(r.get("mesh"), WarpFreeMesh) else WarpFreeMesh(
                r["mesh_vertices"], r["mesh_faces"], cell_size=r.get("cell_size", 0.0))
            st.rc_mesh = self.mesh.device_ptr
        # --- counters
        self.episode_length_buf = z(N, dtype=torch.int64)
        st.episode_length_buf = p(self.episode_length_buf)
        st.max_episode_length = int(s["max_episode_length"])
        st.step_dt, st.physics_dt = float(s["step_dt"]), float(s["physics_dt"])
        st.max_episode_length_s = float(s["max_episode_length_s"])
        # --- user (Python) terms: one row of ext_values per reward / termination term, `dim` rows per observation term
        from . import term_dims as _tdx
        row = 0
        for t in list(s["terminations"]) + list(s["rewards"]):
            if t["op"] == "external":
                t["ext_row"], row = row, row + 1
        for t in s["observations"]["terms"]:
            if t["op"] == "external":
                t["ext_row"], row = row, row + _tdx.term_dim(s, t)
        self.n_ext = row
        self.ext_values = z(max(row, 1), N)
        st.ext_values = p(self.ext_values) if row else None
        # --- terminations / rewards
        nt, nr = len(s["terminations"]), len(s["rewards"])
        if max(nt, nr, len(s["observations"]["terms"])) > C["B200LAB_MAX_TERMS"]:
            raise ValueError("too many terms")
        st.num_term, st.num_rew = nt, nr
        for i, t in enumerate(s["terminations"]):
            self._term(st.term[i], "term", t)
        for i, t in enumerate(s["rewards"]):
            self._term(st.rew[i], "rew", t)
        self.term_dones = z(max(nt, 1), N, dtype=torch.uint8)
        # what a caller reads every step lives in ONE slab so a host consumer needs one D2H copy:
        # [obs N*D f32 | reward N f32 | terminated N u8 | truncated N u8 | reset_buf N u8]
        from . import term_dims as _td
        D_all = sum(_td.term_dim(s, t) for t in s["observations"]["terms"])
        self.result = z(4 * N * D_all + 4 * N + 4 * N, dtype=torch.uint8)
        o_rew, o_msk = 4 * N * D_all, 4 * N * D_all + 4 * N
        self.obs_buf = self.result[:o_rew].view(torch.float32).view(N, D_all)
        self.reward_buf = self.result[o_rew:o_msk].view(torch.float32)
        self.terminated, self.truncated, self.reset_buf = (self.result[o_msk + i * N:o_msk + (i + 1) * N] for i in range(3))
        self.result_layout = {"obs": (0, o_rew), "reward": (o_rew, o_msk), "terminated": (o_msk, o_msk + N),
                              "truncated": (o_msk + N, o_msk + 2 * N), "reset_buf": (o_msk + 2 * N, o_msk + 3 * N)}
        self.episode_sums, self.step_reward = z(max(nr, 1), N), z(N, max(nr, 1))
        st.term_dones, st.terminated, st.truncated, st.reset_buf = p(self.term_dones), p(self.terminated), p(self.truncated), p(self.reset_buf)
        st.reward_buf, st.episode_sums, st.step_reward = p(self.reward_buf), p(self.episode_sums), p(self.step_reward)
        # --- command
        c = s.get("command")
        st.has_command = 1 if c else 0
        if c:
            self.cmd_vel_b, self.cmd_heading_target = z(N, 3), z(N)
            self.cmd_is_heading, self.cmd_is_standing = z(N, dtype=torch.uint8), z(N, dtype=torch.uint8)
            self.cmd_time_left, self.cmd_counter = z(N), z(N, dtype=torch.int64)
            self.cmd_metric_xy, self.cmd_metric_yaw = z(N), z(N)
            st.cmd_vel_b, st.cmd_heading_target = p(self.cmd_vel_b), p(self.cmd_heading_target)
            st.cmd_is_heading, st.cmd_is_standing = p(self.cmd_is_heading), p(self.cmd_is_standing)
            st.cmd_time_left, st.cmd_counter = p(self.cmd_time_left), p(self.cmd_counter)
            st.cmd_metric_xy, st.cmd_metric_yaw = p(self.cmd_metric_xy), p(self.cmd_metric_yaw)
            st.cmd_heading_command = 1 if c["heading_command"] else 0
            lo, hi = c["resampling_time_range"]
            st.cmd_resample_lo, st.cmd_resample_span = float(lo), _f32span_tensor(lo, hi)
            rg = c["ranges"]
            for k, key in enumerate(("lin_vel_x", "lin_vel_y", "ang_vel_z", "heading")):
                lo, hi = rg.get(key) or (0.0, 0.0)
                st.cmd_range[2 * k], st.cmd_range[2 * k + 1] = float(lo), float(hi)
                st.cmd_span[k] = _f32span_tensor(lo, hi)
            st.cmd_rel_standing, st.cmd_rel_heading = float(c["rel_standing_envs"]), float(c.get("rel_heading_envs", 1.0))
            st.cmd_heading_stiffness = float(c.get("heading_control_stiffness", 1.0))
            st.cmd_max_command_step = float(c["resampling_time_range"][1] / s["step_dt"])
        # --- events
        ev = s.get("events", {})
        e = ev.get("reset_root_state")
        st.ev_reset_root_state = 1 if e else 0
        if e:
            for k, key in enumerate(_KEYS6):
                lo, hi = e["pose_range"].get(key, (0.0, 0.0))
                st.ev_pose_range[2 * k], st.ev_pose_range[2 * k + 1] = float(lo), _f32span_tensor(lo, hi)
                lo, hi = e["velocity_range"].get(key, (0.0, 0.0))
                st.ev_vel_range[2 * k], st.ev_vel_range[2 * k + 1] = float(lo), _f32span_tensor(lo, hi)
        def gate_state(min_steps):
            """last-triggered step / triggered-once buffers of a gated reset term (event_manager.py:232-270)."""
            if not min_steps:
                return None, None
            last, once = z(N, dtype=torch.int32), z(N, dtype=torch.uint8)
            self._keep += [last, once]
            return last, once

        self.ev_gate_state = {}
        if e:
            st.ev_root_min_steps = int(e.get("min_step_count", 0))
            last, once = gate_state(st.ev_root_min_steps)
            st.ev_root_last_step, st.ev_root_once = p(last), p(once)
            self.ev_gate_state["reset_root_state"] = (last, once)
        rj = ev.get("reset_joints", [])
        if len(rj) > C["B200LAB_MAX_RESET_JOINT_EVENTS"]:
            raise ValueError("too many reset_joints_by_* events")
        st.ev_num_reset_joints = len(rj)
        for k, e in enumerate(rj):
            st.ev_reset_joints_mode[k] = 1 if e["mode"] == "scale" else 2
            st.ev_reset_joints_min_steps[k] = int(e.get("min_step_count", 0))
            last, once = gate_state(st.ev_reset_joints_min_steps[k])
            st.ev_reset_joints_last_step[k], st.ev_reset_joints_once[k] = p(last), p(once)
            self.ev_gate_state[f"reset_joints_{k}"] = (last, once)
            lo, hi = e["position_range"]
            st.ev_joint_pos_range[k][0], st.ev_joint_pos_range[k][1] = float(lo), _f32span_py(lo, hi)
            lo, hi = e["velocity_range"]
            st.ev_joint_vel_range[k][0], st.ev_joint_vel_range[k][1] = float(lo), _f32span_py(lo, hi)
        pushes = ev.get("interval_push", [])
        if len(pushes) > C["B200LAB_MAX_INTERVAL_EVENTS"]:
            raise ValueError("too many interval events")
        st.num_interval = len(pushes)
        self.interval_time_left = []
        self.interval_global_time = {}
        for k, e in enumerate(pushes):
            t = z(N)
            self.interval_time_left.append(t)
            st.ev_interval_time_left[k] = p(t)
            st.ev_interval_global[k] = 1 if e.get("is_global_time") else 0
            if e.get("is_global_time"):   # one timer for all envs, two slots indexed by step parity
                gt = z(2)
                self.interval_global_time[k] = gt
                st.ev_interval_global_time[k] = p(gt)
            lo, hi = e["interval_range_s"]
            st.ev_interval_range[k][0], st.ev_interval_range[k][1] = float(lo), _f32span_py(lo, hi)
            for a, key in enumerate(_KEYS6):
                lo, hi = e["velocity_range"].get(key, (0.0, 0.0))
                st.ev_push_range[k][2 * a], st.ev_push_range[k][2 * a + 1] = float(lo), _f32span_tensor(lo, hi)
        # --- curriculum / origins
        cur = s.get("curriculum")
        self.env_origins = torch.as_tensor(np.asarray(s["env_origins"]), dtype=torch.float32).to(dev).contiguous() \
            if s.get("env_origins") is not None else z(N, 3)
        st.env_origins = p(self.env_origins)
        st.has_terrain_curriculum = 1 if cur else 0
        if cur:
            self.terrain_origins = const(cur["terrain_origins"])
            self.terrain_levels = torch.as_tensor(np.asarray(cur["terrain_levels"]), dtype=torch.int64).to(dev).contiguous()
            self.terrain_types = const(cur["terrain_types"], torch.int64)
            st.terrain_origins, st.terrain_levels, st.terrain_types = p(self.terrain_origins), p(self.terrain_levels), p(self.terrain_types)
            st.terrain_rows, st.terrain_cols = int(self.terrain_origins.shape[0]), int(self.terrain_origins.shape[1])
            st.terrain_size_x = float(cur["terrain_size_x"])
        # --- observations
        o = s["observations"]
        st.num_obs = len(o["terms"])
        st.obs_enable_corruption = 1 if o.get("enable_corruption", True) else 0
        D = 0
        from . import term_dims
        hist_total = 0
        _MOD = {"scale": 1, "bias": 2, "clip": 3}
        for i, t in enumerate(o["terms"]):
            dst = st.obs[i]
            self._term(dst, "obs", t)
            dst.dim = term_dims.term_raw_dim(s, t)
            D += term_dims.term_dim(s, t)
            H = int(t.get("history_length", 0) or 0)
            if H:   # ring [H, dim] inside the env's obs_hist row (observation_manager.py:422-433)
                dst.history_length, dst.hist_offset = H, hist_total
                hist_total += H * dst.dim
            q = 0
            for k, m in enumerate(t.get("modifiers") or ()):
                dst.mod_kind[k] = _MOD[m[0]]
                for v in m[1:]:
                    dst.mod_p[q] = float(v)
                    q += 1
                dst.num_mod = k + 1
            nz = t.get("noise")
            dst.noise_kind = C["B200LAB_NOISE_NONE"]
            if nz is not None:
                kind, lo, hi, mode = nz
                if kind not in ("uniform", "gaussian", "constant"):
                    raise ValueError(f"unknown observation noise kind '{kind}'")
                dst.noise_kind = C[f"B200LAB_NOISE_{kind.upper()}_{mode.upper()}"]
                # uniform: (n_min, n_max - n_min); gaussian: (mean, std); constant: (bias, -)
                dst.noise_lo = float(lo)
                dst.noise_span = _f32span_py(lo, hi) if kind == "uniform" else float(hi)
            dst.clip_on = 1 if t.get("clip") is not None else 0
            if dst.clip_on:
                dst.clip_lo, dst.clip_hi = float(t["clip"][0]), float(t["clip"][1])
            dst.scale_on = 1 if t.get("scale") is not None else 0
            dst.scale = float(t["scale"]) if dst.scale_on else 1.0
        self.obs_hist = None
        self._hist_started = False
        if hist_total:
            self.obs_hist = z(N, hist_total)
            self.obs_hist_count = z(1, dtype=torch.int32)
            self.obs_hist_reset_step = z(N, dtype=torch.int64)   # = the first step counter: the first compute fills every ring
            st.obs_hist_total = hist_total
            st.obs_hist, st.obs_hist_count = p(self.obs_hist), p(self.obs_hist_count)
            st.obs_hist_reset_step = p(self.obs_hist_reset_step)
        st.obs_dim = D
        self.obs_dim = D
        assert D == self.obs_buf.shape[1]
        st.obs_buf = p(self.obs_buf)
        # --- randomness / logs
        st.seed = int(seed)
        self.n_u_reset = (C["B200LAB_U_RESET_JOINT"] + 2 * J * max(1, len(rj))
                          + sum(1 for a_ in s["actuators"] if a_["kind"] in ("delayed_pd", "remotized_pd")))
        self.n_u_step = C["B200LAB_U_STEP_INTERVAL"] + 7 * len(pushes)
        st.n_u_reset, st.n_u_step = self.n_u_reset, self.n_u_step
        self.log_accum = z(C["B200LAB_LOG_SLOTS"], dtype=torch.float64)
        st.log_accum = p(self.log_accum)
        self.log_running = z(2 * C["B200LAB_LOG_SLOTS"], dtype=torch.float64)
        st.log_running = p(self.log_running)
        self.any_reset = z(1, dtype=torch.int32)
        st.any_reset_dev = p(self.any_reset)
        self._u = (None, None, None)
        self._view_arrays = {}
        self.handle = self.L.b200lab_env_create(ctypes.byref(st))
        if not self.handle:
            raise RuntimeError("b200lab_env_create failed: " + self.L.b200lab_last_error().decode())
        self.step_counter = 0
        self.speculative_scan = False   # opt-in: see b200lab_height_scan_prefetch

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self.L.b200lab_env_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    def interval_timer(self, k):
        """time_left of interval term k as the reference holds it (event_manager.py:368-372): [N] per env, or [1] for an
        `is_global_time` term -- whose two device slots alternate by step parity, so that env 0 can write the new value
        while the other envs of the same launch still read the old one."""
        gt = self.interval_global_time.get(k)
        if gt is None:
            return self.interval_time_left[k]
        t = getattr(self, "step_counter_t", None)
        step = int(t.item()) + 1 if t is not None else self.step_counter
        return gt[step & 1].reshape(1)

    def use_device_step_counter(self):
        """Keeps the step counter (the Philox key of perf mode) on the device: `process_action` / `decimation_fused`
        increment it, the other kernels read it, and the `step_counter` argument of the entry points is ignored -- which
        is what makes a whole step capturable in a CUDA graph."""
        if getattr(self, "step_counter_t", None) is None:
            self.step_counter_t = torch.full((1,), self.step_counter - 1, dtype=torch.int64, device=self.device)
            self._st.step_counter_dev = _lib.ptr(self.step_counter_t)
            _lib.check(self.L.b200lab_env_update(self.handle, ctypes.byref(self._st)))

    # ------------------------------------------------------------------------------------------------
    def set_uniforms(self, u_reset=None, u_step=None, u_obs=None):
        """Parity mode: pre-drawn uniforms in [0,1) replace the in-kernel Philox streams (None -> Philox)."""
        st, p = self._st, _lib.ptr
        for name, t, width in (("u_reset", u_reset, self.n_u_reset), ("u_step", u_step, self.n_u_step),
                               ("u_obs", u_obs, self.obs_dim)):
            if t is not None and tuple(t.shape) != (self.N, width):
                raise ValueError(f"{name} must have shape ({self.N}, {width}), got {tuple(t.shape)}")
        self._u = (u_reset, u_step, u_obs)
        st.u_reset, st.u_step, st.u_obs = p(u_reset), p(u_step), p(u_obs)
        _lib.check(self.L.b200lab_env_update(self.handle, ctypes.byref(st)))

    def process_action(self, action):
        if action.shape != (self.N, self.A):
            raise ValueError(f"Invalid action shape, expected: {self.A}, received: {action.shape[1]}.")
        _lib.check(self.L.b200lab_process_action(self.handle, _lib.ptr(action.contiguous()), _lib.current_stream()))

    def substep_pre(self, sim: SimView):
        _lib.check(self.L.b200lab_substep_pre(self.handle, sim.ref, _lib.current_stream()))

    def substep_post(self, sim: SimView):
        _lib.check(self.L.b200lab_substep_post(self.handle, sim.ref, _lib.current_stream()))

    def height_scan_prefetch(self, sim: SimView):
        """Start the height scan from the post-physics state `sim` on the library's side stream; the next `compute_obs`
        joins it (b200lab_height_scan_prefetch).  `step` does this when `speculative_scan` is set."""
        _lib.check(self.L.b200lab_height_scan_prefetch(self.handle, sim.ref, _lib.current_stream()))

    def post_step(self, sim: SimView):
        _lib.check(self.L.b200lab_post_step(self.handle, sim.ref, self.step_counter, _lib.current_stream()))

    def log_fold(self):
        """Folds the per-step log sums into extras["log"] values (b200lab_log_fold); `step_tail` does it itself."""
        _lib.check(self.L.b200lab_log_fold(self.handle, _lib.current_stream()))

    def log_keys(self):
        """Keys of extras["log"] in slot order 1.. (manager_based_rl_env.py:365-389 of the reference)."""
        s = self.spec
        keys = ["Episode_Reward/" + t["name"] for t in s["rewards"]]
        keys += ["Episode_Termination/" + t["name"] for t in s["terminations"]]
        cname = s["command"]["name"] if s.get("command") else "base_velocity"
        keys += [f"Metrics/{cname}/error_vel_xy", f"Metrics/{cname}/error_vel_yaw"]
        keys += ["Curriculum/" + (s["curriculum"]["name"] if s.get("curriculum") else "terrain_levels")]
        return keys

    def _log_dict(self, vals):
        s = self.spec
        out = {}
        for i, k in enumerate(self.log_keys()):
            if k.startswith("Metrics/") and not s.get("command"):
                continue
            if k.startswith("Curriculum/") and not s.get("curriculum"):
                continue
            out[k] = float(vals[1 + i])
        return out

    def last_step_log(self):
        """extras["log"] of the most recent step (one device->host read; {} when no env was reset in it)."""
        S = C["B200LAB_LOG_SLOTS"]
        v = self.log_running[S:].cpu()
        return self._log_dict(v) if float(v[0]) > 0 else {}

    def mean_log(self, reset: bool = True):
        """Mean of the per-step extras["log"] dicts since the last call -- what the reference runner computes from the
        dicts it collects (on_policy_runner.py:log) -- from ONE device->host read."""
        S = C["B200LAB_LOG_SLOTS"]
        v = self.log_running[:S].cpu()
        if reset:
            self.log_running[:S].zero_()
        n = float(v[0])
        return self._log_dict(v / n) if n > 0 else {}

    def post_reset_write(self, sim: SimView):
        """`scene.write_data_to_sim()` after `_reset_idx` (manager_based_rl_env.py:217-223); device-side condition."""
        _lib.check(self.L.b200lab_post_reset_write(self.handle, sim.ref, _lib.current_stream()))

    def step_tail(self, sim: SimView):
        """post_reset_write + compute_obs, overlapped (b200lab_step_tail)."""
        _lib.check(self.L.b200lab_step_tail(self.handle, sim.ref, self.step_counter, _lib.current_stream()))
        return self.obs_buf

    def compute_obs(self, sim: SimView, in_step: bool = False):
        """ObservationManager.compute.  Outside a step (the observation `env.reset()` returns) the history rings of every
        env are filled by the first call, and no env is re-filled by the next step's compute unless that step resets it
        (b200lab.h: obs_hist_reset_step)."""
        standalone = self.obs_hist is not None and not in_step
        if standalone and not self._hist_started:
            t = getattr(self, "step_counter_t", None)
            if t is not None:
                self.obs_hist_reset_step.copy_(t.expand(self.N))
            else:
                self.obs_hist_reset_step.fill_(self.step_counter)
        _lib.check(self.L.b200lab_compute_obs(self.handle, sim.ref, self.step_counter, _lib.current_stream()))
        if standalone:
            self.obs_hist_reset_step.fill_(-1)
        self._hist_started = True
        return self.obs_buf

    def step_tail_ranges(self, sim: SimView, ranges, after_range):
        """`step_tail` with the observation rows computed range by range (b200lab_step_tail_begin /
        b200lab_compute_obs_range / b200lab_step_tail_end); `after_range(begin, end)` runs after each range has been
        queued, e.g. to start its device->host copy on another stream while the next range is computed."""
        st = _lib.current_stream()
        _lib.check(self.L.b200lab_step_tail_begin(self.handle, sim.ref, st))
        for b, e in ranges:
            _lib.check(self.L.b200lab_compute_obs_range(self.handle, sim.ref, self.step_counter, b, e, st))
            after_range(b, e)
        _lib.check(self.L.b200lab_step_tail_end(self.handle, st))
        return self.obs_buf

    def decimation_fused(self, action, sims, sim_prev):
        """Replay fast path: process_action + all S (pre, post) sub-steps in one launch (b200lab_decimation_fused)."""
        if action.shape != (self.N, self.A):
            raise ValueError(f"Invalid action shape, expected: {self.A}, received: {action.shape[1]}.")
        key = tuple(id(v) for v in sims)
        arr = self._view_arrays.get(key)
        if arr is None:
            if len(self._view_arrays) > 256:
                self._view_arrays.clear()
            arr = (_lib.structs["B200LabSimView"] * len(sims))(*[v.c for v in sims])
            self._view_arrays[key] = arr
        action = action.contiguous()
        _lib.check(self.L.b200lab_decimation_fused(self.handle, _lib.ptr(action), sim_prev.ref, arr, len(sims),
                                                   _lib.current_stream()))

    def step(self, action, sims, sim_prev, fused: bool = True, obs_ranges=None, after_post_step=None, after_range=None,
             before_post_step=None):
        """One env step on S known post-physics states: [process_action -> S x (pre(previous state), post(new state))]
        -> post_step -> compute_obs (manager_based_rl_env.py:153-242; mirrors oracle.mdp_env.OracleEnv.step).
        `fused=True` runs the bracketed part as one launch, `fused=False` as the 1 + 2S calls a live simulator uses."""
        if fused and len(sims) <= 8:
            self.decimation_fused(action, sims, sim_prev)
            prev = sims[-1]
        else:
            self.process_action(action)
            prev = sim_prev
            for sim in sims:
                self.substep_pre(prev)
                self.substep_post(sim)
                prev = sim
        if self.speculative_scan:
            self.height_scan_prefetch(prev)
        if before_post_step is not None:   # user reward / termination terms are evaluated on the post-physics state
            before_post_step()
        self.post_step(prev)
        if after_post_step is not None:
            after_post_step()
        if fused and obs_ranges is not None:
            self.step_tail_ranges(prev, obs_ranges, after_range)
        elif fused:
            self.step_tail(prev)
        else:
            self.post_reset_write(prev)
            self.log_fold()
            self.compute_obs(prev, in_step=True)
        self._hist_started = True
        self.step_counter += 1
        return self.obs_buf, self.reward_buf, self.terminated, self.truncated
