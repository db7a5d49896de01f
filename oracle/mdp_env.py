"""Oracle restatement of the non-physics part of ManagerBasedRLEnv.step in plain torch on CPU.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Paths: LAB = source/isaaclab/isaaclab,
TASKS = source/isaaclab_tasks/isaaclab_tasks of the reference.

The environment is described by a plain-dict `spec` (the same neutral description the product compiles its
C structs from; documented in DESIGN.md "Env spec").  PhysX is not restated: its outputs enter as `sim` dicts
(root_state_w [N,13], joint_pos [N,J], joint_vel [N,J], net_forces_w [N,B,3], body_lin_vel_w [N,B,3]) -- one per
physics sub-step.  Every random draw of the reference is made explicit as a pre-drawn uniform in [0,1):
u_reset [N, n_u_reset], u_step [N, n_u_step], u_obs [N, obs_dim]; slot layout = include/b200lab.h B200LAB_U_*.
"""
from __future__ import annotations

import math

import torch

from . import math_utils as mu
from . import raycast as rc

U_RESET_CURRICULUM, U_RESET_DRIFT, U_RESET_POSE, U_RESET_VEL, U_RESET_CMD, U_RESET_JOINT = 0, 1, 4, 10, 16, 23
U_STEP_CMD, U_STEP_INTERVAL = 0, 7

F32 = torch.float32


def _t(x, dtype=F32):
    return x.clone().to(dtype) if isinstance(x, torch.Tensor) else torch.tensor(x, dtype=dtype)


DELAYED_KINDS = ("delayed_pd", "remotized_pd")


def linear_interpolation(x, y, q):
    """LAB/utils/interpolation/linear_interpolation.py:53-77: piecewise-linear y(x) at q, zero-order hold outside."""
    q1 = q.reshape(-1)
    n_smaller = torch.sum(x.unsqueeze(1) < q1.unsqueeze(0), dim=0, dtype=torch.int64)
    lo = torch.clamp(n_smaller - 1, min=0)
    hi = torch.clamp(n_smaller, max=x.numel() - 1)
    w = (q1 - x[lo]) / (x[hi] - x[lo])
    w[hi == lo] = 0.0
    return (y[lo] + w * (y[hi] - y[lo])).reshape(q.shape)


def u_reset_delay(spec):
    """First of the per-group time-lag slots of the DelayedPD actuator groups (one draw per group and reset)."""
    return U_RESET_JOINT + 2 * spec["J"] * max(1, len(spec.get("events", {}).get("reset_joints", [])))


def n_u_reset(spec):
    return u_reset_delay(spec) + sum(1 for a in spec["actuators"] if a["kind"] in DELAYED_KINDS)


def n_u_step(spec):
    return U_STEP_INTERVAL + 7 * len(spec.get("events", {}).get("interval_push", []))


def obs_dim(spec):
    return sum(term_dim(spec, t) for t in spec["observations"]["terms"])


def term_dim(spec, t):
    """Width of the term in the concatenated group: its own width times the flattened history length
    (observation_manager.py:422-433)."""
    return term_raw_dim(spec, t) * max(1, int(t.get("history_length", 0) or 0))


def term_raw_dim(spec, t):
    op = t["op"]
    if op in ("base_lin_vel", "base_ang_vel", "projected_gravity", "generated_commands"):
        return 3
    if op in ("joint_pos_rel", "joint_vel_rel", "joint_pos", "joint_vel"):
        return len(t.get("joint_ids", range(spec["J"])))
    if op == "last_action":
        return spec["A"]
    if op == "height_scan":
        return spec["R"]
    if op == "base_pos_z":
        return 1
    raise ValueError(op)


def apply_noise(x, u, nz):
    """`nz = (kind, a, b, operation)`; `u` = the element-wise random draw the reference function makes (rand_like for
    uniform, randn_like for gaussian, unused for constant)."""
    kind, lo, hi, mode = nz
    if kind == "uniform":       # noise_model.py:45-70; (lo, hi) = (n_min, n_max)
        if mode == "add":
            return x + u * (hi - lo) + lo
        if mode == "scale":
            return x * (u * (hi - lo) + lo)
        return u * (hi - lo) + lo
    if kind == "gaussian":      # noise_model.py:73-98; (lo, hi) = (mean, std)
        if mode == "add":
            return x + lo + hi * u
        if mode == "scale":
            return x * (lo + hi * u)
        return lo + hi * u
    assert kind == "constant"   # noise_model.py:20-42; lo = bias
    if mode == "add":
        return x + lo
    if mode == "scale":
        return x * lo
    return torch.zeros_like(x) + lo


class Ring:
    """Batched ring buffer with the semantics of LAB/utils/buffers/circular_buffer.py:13-160: one write pointer for all
    envs; the first push after a reset fills every slot of that env with the pushed row."""

    def __init__(self, length, N):
        self.L, self.N = int(length), N
        self.buf, self.ptr = None, -1
        self.pushes = torch.zeros(N, dtype=torch.int64)

    def reset(self, ids):
        self.pushes[ids] = 0
        if self.buf is not None:
            self.buf[:, ids] = 0.0

    def append(self, x):
        if self.buf is None:
            self.buf = torch.zeros(self.L, *x.shape, dtype=x.dtype)
        self.ptr = (self.ptr + 1) % self.L
        self.buf[self.ptr] = x
        fresh = self.pushes == 0
        self.buf[:, fresh] = x[fresh]
        self.pushes += 1

    def lagged(self, lags):
        """Row pushed `lags` appends ago, per env (capped at what has been pushed): circular_buffer.py:140-160."""
        k = torch.minimum(lags.to(torch.int64), self.pushes - 1)
        return self.buf[torch.remainder(self.ptr - k, self.L), torch.arange(self.N)]

    def oldest_first(self):
        """[N, L, ...] with the newest row last (circular_buffer.py:74-83)."""
        return torch.roll(self.buf, shifts=self.L - self.ptr - 1, dims=0).transpose(0, 1)


def apply_modifiers(x, mods):
    """Function modifiers of an observation term, in order (LAB/utils/modifiers/modifier.py:23-66)."""
    for m in mods or ():
        if m[0] == "scale":
            x = x * m[1]
        elif m[0] == "bias":
            x = x + m[1]
        elif m[0] == "clip":
            x = x.clip(min=m[1], max=m[2])
        else:
            raise ValueError(m[0])
    return x


class OracleEnv:
    def __init__(self, spec):
        self.spec = s = spec
        N, J, A, B = s["N"], s["J"], s["A"], s["B"]
        self.N, self.J, self.A, self.B, self.R = N, J, A, B, s.get("R", 0)
        z = lambda *shape: torch.zeros(*shape, dtype=F32)  # noqa: E731
        # articulation buffers (LAB/assets/articulation/articulation.py:1207-1266)
        self.default_root_state = _t(s["default_root_state"])
        self.default_joint_pos = _t(s["default_joint_pos"])
        self.default_joint_vel = _t(s["default_joint_vel"])
        self.soft_joint_pos_limits = _t(s["soft_joint_pos_limits"])
        self.soft_joint_vel_limits = _t(s["soft_joint_vel_limits"])
        self.joint_pos_target, self.joint_vel_target, self.joint_effort_target = z(N, J), z(N, J), z(N, J)
        self.computed_torque, self.applied_torque = z(N, J), z(N, J)
        self.joint_pos_target_sim, self.joint_effort_target_sim = z(N, J), z(N, J)
        self.prev_joint_vel, self.joint_acc = z(N, J), z(N, J)
        # action manager
        self.action, self.prev_action, self.raw_actions, self.processed_actions = z(N, A), z(N, A), z(N, A), z(N, A)
        # actuators
        self.act_state = []
        for a in s["actuators"]:
            Jg = len(a["joint_ids"])
            st = {}
            if a["kind"] == "lstm":
                L, H = a["lstm"]["num_layers"], a["lstm"]["hidden"]
                st["h"], st["c"] = z(L, N * Jg, H), z(L, N * Jg, H)
            elif a["kind"] == "mlp":
                hist = max(a["mlp"]["input_idx"]) + 1
                st["pos_hist"], st["vel_hist"] = z(N, hist, Jg), z(N, hist, Jg)
            elif a["kind"] in DELAYED_KINDS:   # DelayBuffer x3 with shared lags (actuator_pd.py:303-330, delay_buffer.py)
                st["ring"] = Ring(a["max_delay"] + 1, N)
                st["lags"] = torch.zeros(N, dtype=torch.int64)
            self.act_state.append(st)
        # observation history (observation_manager.py:384-433)
        self.obs_hist = {i: Ring(t["history_length"], N) for i, t in enumerate(s["observations"]["terms"])
                         if t.get("history_length")}
        # contact sensor (LAB/sensors/contact_sensor/contact_sensor.py)
        if s.get("contact"):
            H = s["contact"]["history_length"]
            self.cs_net_forces_w, self.cs_hist = z(N, B, 3), z(N, H, B, 3)
            self.cs_last_air, self.cs_cur_air, self.cs_last_contact, self.cs_cur_contact = z(N, B), z(N, B), z(N, B), z(N, B)
            self.cs_ts, self.cs_ts_last, self.cs_outdated = z(N), z(N), torch.ones(N, dtype=torch.bool)
        # ray caster (LAB/sensors/ray_caster/ray_caster.py)
        if s.get("raycaster"):
            R = self.R
            self.rc_ts, self.rc_ts_last, self.rc_outdated = z(N), z(N), torch.ones(N, dtype=torch.bool)
            self.rc_drift, self.rc_pos_w, self.rc_quat_w, self.rc_hits = z(N, 3), z(N, 3), z(N, 4), z(N, R, 3)
            self.mesh = rc.Mesh(s["raycaster"]["mesh_vertices"], s["raycaster"]["mesh_faces"])
        # env counters
        self.episode_length_buf = torch.zeros(N, dtype=torch.int64)
        # managers
        nt, nr = len(s["terminations"]), len(s["rewards"])
        self.term_dones = torch.zeros(nt, N, dtype=torch.bool)
        self.terminated, self.truncated = torch.zeros(N, dtype=torch.bool), torch.zeros(N, dtype=torch.bool)
        self.reset_buf = torch.zeros(N, dtype=torch.bool)
        self.reward_buf, self.episode_sums, self.step_reward = z(N), z(nr, N), z(N, nr)
        if s.get("command"):
            self.cmd_vel_b, self.cmd_heading_target = z(N, 3), z(N)
            self.cmd_is_heading = torch.zeros(N, dtype=torch.bool)
            self.cmd_is_standing = torch.zeros(N, dtype=torch.bool)
            self.cmd_time_left, self.cmd_counter = z(N), torch.zeros(N, dtype=torch.int64)
            self.cmd_metric_xy, self.cmd_metric_yaw = z(N), z(N)
        ev = s.get("events", {})
        self.interval_time_left = [z(N) for _ in ev.get("interval_push", [])]
        # is_global_time terms keep ONE timer (event_manager.py:113-118); gated reset terms keep the last trigger step
        self.interval_global_time = {k: torch.zeros(1) for k, e in enumerate(ev.get("interval_push", [])) if e.get("is_global_time")}
        self.common_step_counter = 0
        self._gate = {}
        cur = s.get("curriculum")
        if cur:
            self.terrain_origins = _t(cur["terrain_origins"])
            self.terrain_levels = _t(cur["terrain_levels"], torch.int64)
            self.terrain_types = _t(cur["terrain_types"], torch.int64)
        self.env_origins = _t(s["env_origins"]) if s.get("env_origins") is not None else z(N, 3)
        self.obs_buf = z(N, obs_dim(s))
        self.log = {}

    # ---------------------------------------------------------------------------------------------
    # derived articulation state (LAB/assets/articulation/articulation_data.py:513-526,604-619)
    # ---------------------------------------------------------------------------------------------
    @staticmethod
    def derived(sim):
        rs = sim["root_state_w"]
        q = rs[:, 3:7]
        d = dict(pos=rs[:, 0:3], quat=q, lin_w=rs[:, 7:10], ang_w=rs[:, 10:13])
        d["lin_b"] = mu.quat_rotate_inverse(q, rs[:, 7:10])
        d["ang_b"] = mu.quat_rotate_inverse(q, rs[:, 10:13])
        g = torch.tensor([0.0, 0.0, -1.0]).repeat(rs.shape[0], 1)  # GRAVITY_VEC_W (articulation.py:1275-1277)
        d["grav_b"] = mu.quat_rotate_inverse(q, g)
        fwd = mu.quat_apply(q, torch.tensor([1.0, 0.0, 0.0]).repeat(rs.shape[0], 1))
        d["heading_w"] = torch.atan2(fwd[:, 1], fwd[:, 0])
        return d

    # ---------------------------------------------------------------------------------------------
    # ActionManager.process_action + JointAction.process_actions
    # (LAB/managers/action_manager.py:318-339, LAB/envs/mdp/actions/joint_actions.py:130-139)
    # ---------------------------------------------------------------------------------------------
    def process_action(self, action):
        a = self.spec["action"]
        self.prev_action[:] = self.action
        self.action[:] = action
        self.raw_actions[:] = action
        self.processed_actions = self.raw_actions * _t(a["scale"]) + _t(a["offset"])
        if a.get("clip") is not None:
            c = _t(a["clip"])
            self.processed_actions = torch.clamp(self.processed_actions, min=c[:, 0], max=c[:, 1])

    # ---------------------------------------------------------------------------------------------
    # apply_action + Articulation._apply_actuator_model (articulation.py:1432-1469) with the actuator models
    # of LAB/actuators/actuator_pd.py:115-140,184-198,264-286 and actuator_net.py:76-99,147-189
    # ---------------------------------------------------------------------------------------------
    def substep_pre(self, sim):
        self.apply_action()
        self.write_data_to_sim(sim)

    def apply_action(self):
        """ActionManager.apply_action -> JointPositionAction/JointEffortAction.apply_actions (joint_actions.py:163-186)."""
        a = self.spec["action"]
        ids = list(a["joint_ids"])
        if a["kind"] == "joint_position":
            self.joint_pos_target[:, ids] = self.processed_actions
        else:
            self.joint_effort_target[:, ids] = self.processed_actions

    def write_data_to_sim(self, sim):
        """Articulation.write_data_to_sim (articulation.py:173-200): the actuator models advance on EVERY call -- also on
        the one `step` makes after `_reset_idx` (manager_based_rl_env.py:222), for all envs, with the targets of the
        last `apply_action` and the joint state that holds the reset writes."""
        s = self.spec
        for act, st in zip(s["actuators"], self.act_state):
            j = list(act["joint_ids"])
            q, qd = sim["joint_pos"][:, j], sim["joint_vel"][:, j]
            q_t, qd_t, ff = self.joint_pos_target[:, j], self.joint_vel_target[:, j], self.joint_effort_target[:, j]
            elim = _t(act["effort_limit"]).expand(len(j))
            kind = act["kind"]
            if kind in DELAYED_KINDS:   # the three set-points go through the delay line, then IdealPD (actuator_pd.py:336-347)
                st["ring"].append(torch.stack([q_t, qd_t, ff], dim=-1))
                d3 = st["ring"].lagged(st["lags"])
                q_t, qd_t, ff = d3[..., 0], d3[..., 1], d3[..., 2]
            if kind in ("implicit", "ideal_pd", "dc_motor") + DELAYED_KINDS:
                err_p, err_v = q_t - q, qd_t - qd
                computed = _t(act["stiffness"]) * err_p + _t(act["damping"]) * err_v + ff
            elif kind == "lstm":
                computed = self._lstm(act["lstm"], st, q_t - q, qd)
            elif kind == "mlp":
                computed = self._mlp(act["mlp"], st, q_t - q, qd)
            else:
                raise ValueError(kind)
            if kind in ("implicit", "ideal_pd") + DELAYED_KINDS:
                applied = torch.clip(computed, min=-elim, max=elim)  # actuator_base.py:309-318
                if kind == "remotized_pd":   # angle-dependent limit at the CURRENT joint position (actuator_pd.py:405-416)
                    lim = linear_interpolation(_t(act["lookup_angle"]), _t(act["lookup_torque"]), q)
                    applied = torch.clamp(applied, min=-lim, max=lim)
            else:  # DCMotor._clip_effort (actuator_pd.py:276-286)
                sat = float(act["saturation_effort"])
                vlim = _t(act["velocity_limit"]).expand(len(j))
                max_e = torch.clip(sat * (1.0 - qd / vlim), min=torch.zeros_like(elim), max=elim)
                min_e = torch.clip(sat * (-1.0 - qd / vlim), min=-elim, max=torch.zeros_like(elim))
                applied = torch.clip(computed, min=min_e, max=max_e)
            self.computed_torque[:, j] = computed
            self.applied_torque[:, j] = applied
            if kind == "implicit":  # control action returned unchanged (actuator_pd.py:140)
                self.joint_pos_target_sim[:, j] = q_t
                self.joint_effort_target_sim[:, j] = ff
            else:
                self.joint_effort_target_sim[:, j] = applied

    @staticmethod
    def _lstm(net, st, pos_err, vel):
        """actuator_net.py:76-99: LSTM(input 2, seq len 1) + Linear head on rows env*Jg+j; state [L, N*Jg, H]."""
        x = torch.stack([pos_err.flatten(), vel.flatten()], dim=-1)
        L = net["num_layers"]
        for l in range(L):
            gates = (x @ net[f"weight_ih_l{l}"].T + net[f"bias_ih_l{l}"] + st["h"][l] @ net[f"weight_hh_l{l}"].T
                     + net[f"bias_hh_l{l}"])
            i, f, g, o = gates.chunk(4, dim=-1)
            c = torch.sigmoid(f) * st["c"][l] + torch.sigmoid(i) * torch.tanh(g)
            h = torch.sigmoid(o) * torch.tanh(c)
            st["h"][l], st["c"][l] = h, c
            x = h
        out = x @ net["head_weight"].T + net["head_bias"]
        return out.reshape(pos_err.shape)

    @staticmethod
    def _mlp(net, st, pos_err, vel):
        """actuator_net.py:147-189."""
        st["pos_hist"] = st["pos_hist"].roll(1, 1)
        st["pos_hist"][:, 0] = pos_err
        st["vel_hist"] = st["vel_hist"].roll(1, 1)
        st["vel_hist"][:, 0] = vel
        N, Jg = pos_err.shape
        pos_in = torch.cat([st["pos_hist"][:, i].unsqueeze(2) for i in net["input_idx"]], dim=2).view(N * Jg, -1)
        vel_in = torch.cat([st["vel_hist"][:, i].unsqueeze(2) for i in net["input_idx"]], dim=2).view(N * Jg, -1)
        if net["input_order"] == "pos_vel":
            x = torch.cat([pos_in * net["pos_scale"], vel_in * net["vel_scale"]], dim=1)
        else:
            x = torch.cat([vel_in * net["vel_scale"], pos_in * net["pos_scale"]], dim=1)
        acts = {"softsign": torch.nn.functional.softsign, "elu": torch.nn.functional.elu,
                "relu": torch.relu, "tanh": torch.tanh}
        n = len(net["layers"])
        for k, (W, b) in enumerate(net["layers"]):
            x = x @ W.T + b
            if k < n - 1:
                x = acts[net["activation"]](x)
        return x.view(N, Jg) * net["torque_scale"]

    # ---------------------------------------------------------------------------------------------
    # scene.update(dt): ArticulationData.update (articulation_data.py:78-83,547-556), SensorBase.update
    # (LAB/sensors/sensor_base.py:197-205,287-296), ContactSensor._update_buffers_impl (contact_sensor.py:320-379)
    # ---------------------------------------------------------------------------------------------
    def substep_post(self, sim):
        s = self.spec
        dt = s["physics_dt"]
        qd = sim["joint_vel"]
        self.joint_acc = (qd - self.prev_joint_vel) / dt
        self.prev_joint_vel[:] = qd
        cs = s.get("contact")
        if cs:
            self.cs_ts += dt
            self.cs_outdated |= self.cs_ts - self.cs_ts_last + 1e-6 >= cs["update_period"]
            ids = self.cs_outdated.nonzero().squeeze(-1)
            if len(ids) > 0:
                self.cs_net_forces_w[ids] = sim["net_forces_w"][ids]
                self.cs_hist[ids, 1:] = self.cs_hist[ids, :-1].clone()
                self.cs_hist[ids, 0] = self.cs_net_forces_w[ids]
                if cs.get("track_air_time", True):
                    elapsed = (self.cs_ts[ids] - self.cs_ts_last[ids]).unsqueeze(-1)
                    is_contact = torch.norm(self.cs_net_forces_w[ids], dim=-1) > cs["force_threshold"]
                    first_contact = (self.cs_cur_air[ids] > 0) * is_contact
                    first_detached = (self.cs_cur_contact[ids] > 0) * ~is_contact
                    self.cs_last_air[ids] = torch.where(first_contact, self.cs_cur_air[ids] + elapsed, self.cs_last_air[ids])
                    self.cs_cur_air[ids] = torch.where(~is_contact, self.cs_cur_air[ids] + elapsed, 0.0)
                    self.cs_last_contact[ids] = torch.where(first_detached, self.cs_cur_contact[ids] + elapsed,
                                                            self.cs_last_contact[ids])
                    self.cs_cur_contact[ids] = torch.where(is_contact, self.cs_cur_contact[ids] + elapsed, 0.0)
                self.cs_ts_last[ids] = self.cs_ts[ids]
                self.cs_outdated[ids] = False
        r = s.get("raycaster")
        if r:  # history_length == 0: only marked outdated, recomputed lazily on .data (sensor_base.py:204)
            self.rc_ts += dt
            self.rc_outdated |= self.rc_ts - self.rc_ts_last + 1e-6 >= r["update_period"]

    # ---------------------------------------------------------------------------------------------
    # term library
    # ---------------------------------------------------------------------------------------------
    def _hist_max_norm(self, body_ids):
        """max over history of |F| for the selected bodies: net_forces_w_history[:, :, ids].norm(-1).max(1)."""
        return torch.max(torch.norm(self.cs_hist[:, :, body_ids], dim=-1), dim=1)[0]

    def _joint_ids(self, t):
        return list(t["joint_ids"]) if t.get("joint_ids") is not None else list(range(self.J))

    def eval_termination(self, t, sim, d):
        op = t["op"]
        if op == "time_out":  # LAB/envs/mdp/terminations.py:30-32
            return self.episode_length_buf >= self.spec["max_episode_length"]
        if op == "illegal_contact":  # terminations.py:150-158
            return torch.any(self._hist_max_norm(list(t["body_ids"])) > t["threshold"], dim=1)
        if op == "joint_pos_out_of_manual_limit":  # terminations.py:90-105
            j = self._joint_ids(t)
            q = sim["joint_pos"][:, j]
            return torch.any(q > t["bounds"][1], dim=1) | torch.any(q < t["bounds"][0], dim=1)
        if op == "root_height_below_minimum":  # terminations.py:74-82
            return d["pos"][:, 2] < t["minimum_height"]
        if op == "bad_orientation":  # terminations.py:60-71
            return torch.acos(-d["grav_b"][:, 2]).abs() > t["limit_angle"]
        raise ValueError(op)

    def eval_reward(self, t, sim, d):
        op = t["op"]
        cmd = self.cmd_vel_b if self.spec.get("command") else None
        if op == "is_alive":  # LAB/envs/mdp/rewards.py:31-33
            return (~self.terminated).float()
        if op == "is_terminated":  # rewards.py:36-38
            return self.terminated.float()
        if op == "lin_vel_z_l2":  # rewards.py:76-80
            return torch.square(d["lin_b"][:, 2])
        if op == "ang_vel_xy_l2":  # rewards.py:83-87
            return torch.sum(torch.square(d["ang_b"][:, :2]), dim=1)
        if op == "flat_orientation_l2":  # rewards.py:90-97
            return torch.sum(torch.square(d["grav_b"][:, :2]), dim=1)
        j = self._joint_ids(t)
        if op == "joint_torques_l2":  # rewards.py:136-143
            return torch.sum(torch.square(self.applied_torque[:, j]), dim=1)
        if op == "joint_vel_l1":  # rewards.py:146-150
            return torch.sum(torch.abs(sim["joint_vel"][:, j]), dim=1)
        if op == "joint_vel_l2":  # rewards.py:153-160
            return torch.sum(torch.square(sim["joint_vel"][:, j]), dim=1)
        if op == "joint_acc_l2":  # rewards.py:163-170
            return torch.sum(torch.square(self.joint_acc[:, j]), dim=1)
        if op == "joint_deviation_l1":  # rewards.py:173-179
            return torch.sum(torch.abs(sim["joint_pos"][:, j] - self.default_joint_pos[j]), dim=1)
        if op == "joint_pos_limits":  # rewards.py:182-196
            q = sim["joint_pos"][:, j]
            out = -(q - self.soft_joint_pos_limits[j, 0]).clip(max=0.0)
            out += (q - self.soft_joint_pos_limits[j, 1]).clip(min=0.0)
            return torch.sum(out, dim=1)
        if op == "joint_pos_target_l2":  # TASKS/manager_based/classic/cartpole/mdp/rewards.py:19-26
            return torch.sum(torch.square(mu.wrap_to_pi(sim["joint_pos"][:, j]) - t["target"]), dim=1)
        if op == "action_rate_l2":  # rewards.py:245-247
            return torch.sum(torch.square(self.action - self.prev_action), dim=1)
        if op == "action_l2":  # rewards.py:250-252
            return torch.sum(torch.square(self.action), dim=1)
        if op == "undesired_contacts":  # rewards.py:260-268
            return torch.sum(self._hist_max_norm(list(t["body_ids"])) > t["threshold"], dim=1).float()
        if op == "contact_forces":  # rewards.py:271-279
            return torch.sum((self._hist_max_norm(list(t["body_ids"])) - t["threshold"]).clip(min=0.0), dim=1)
        if op == "track_lin_vel_xy_exp":  # rewards.py:287-298
            err = torch.sum(torch.square(cmd[:, :2] - d["lin_b"][:, :2]), dim=1)
            return torch.exp(-err / t["std"] ** 2)
        if op == "track_ang_vel_z_exp":  # rewards.py:301-309
            return torch.exp(-torch.square(cmd[:, 2] - d["ang_b"][:, 2]) / t["std"] ** 2)
        # TASKS/manager_based/locomotion/velocity/mdp/rewards.py
        if op == "feet_air_time":  # :25-44
            b = list(t["body_ids"])
            cc = self.cs_cur_contact[:, b]
            first = (cc > 0.0) * (cc < (self.spec["step_dt"] + 1.0e-8))  # compute_first_contact (contact_sensor.py:176-210)
            rew = torch.sum((self.cs_last_air[:, b] - t["threshold"]) * first, dim=1)
            return rew * (torch.norm(cmd[:, :2], dim=1) > 0.1)
        if op == "feet_air_time_positive_biped":  # :47-66
            b = list(t["body_ids"])
            air, con = self.cs_cur_air[:, b], self.cs_cur_contact[:, b]
            in_contact = con > 0.0
            in_mode = torch.where(in_contact, con, air)
            single = torch.sum(in_contact.int(), dim=1) == 1
            rew = torch.min(torch.where(single.unsqueeze(-1), in_mode, 0.0), dim=1)[0]
            rew = torch.clamp(rew, max=t["threshold"])
            return rew * (torch.norm(cmd[:, :2], dim=1) > 0.1)
        if op == "feet_slide":  # :69-83
            contacts = self._hist_max_norm(list(t["body_ids"])) > 1.0
            v = sim["body_lin_vel_w"][:, list(t["asset_body_ids"]), :2]
            return torch.sum(v.norm(dim=-1) * contacts, dim=1)
        if op == "track_lin_vel_xy_yaw_frame_exp":  # :86-96
            v = mu.quat_rotate_inverse(mu.yaw_quat(d["quat"]), d["lin_w"])
            err = torch.sum(torch.square(cmd[:, :2] - v[:, :2]), dim=1)
            return torch.exp(-err / t["std"] ** 2)
        if op == "track_ang_vel_z_world_exp":  # :99-106
            return torch.exp(-torch.square(cmd[:, 2] - d["ang_w"][:, 2]) / t["std"] ** 2)
        raise ValueError(op)

    def eval_obs(self, t, sim, d):
        op = t["op"]
        if op == "base_lin_vel":  # LAB/envs/mdp/observations.py:40-44
            return d["lin_b"]
        if op == "base_ang_vel":  # :47-51
            return d["ang_b"]
        if op == "projected_gravity":  # :54-58
            return d["grav_b"]
        if op == "generated_commands":  # :529-531
            return self.cmd_vel_b
        if op == "base_pos_z":  # :33-37
            return d["pos"][:, 2:3]
        j = self._joint_ids(t)
        if op == "joint_pos_rel":  # :114-121
            return sim["joint_pos"][:, j] - self.default_joint_pos[j]
        if op == "joint_vel_rel":  # :150-157
            return sim["joint_vel"][:, j] - self.default_joint_vel[j]
        if op == "joint_pos":  # :104-111
            return sim["joint_pos"][:, j]
        if op == "joint_vel":  # :140-147
            return sim["joint_vel"][:, j]
        if op == "last_action":  # :512-521
            return self.action
        if op == "height_scan":  # :165-173 ; sensor.data triggers the lazy update (sensor_base.py:287-296)
            self._raycaster_update(sim)
            return self.rc_pos_w[:, 2].unsqueeze(1) - self.rc_hits[..., 2] - t.get("offset", 0.5)
        raise ValueError(op)

    def _raycaster_update(self, sim):
        """RayCaster._update_buffers_impl (ray_caster.py:220-260) + raycast_mesh (LAB/utils/warp/ops.py:24-127)."""
        r = self.spec["raycaster"]
        ids = self.rc_outdated.nonzero().squeeze(-1)
        if len(ids) == 0:
            return
        rs = sim["root_state_w"][ids]
        pos_w = rs[:, 0:3].clone() + self.rc_drift[ids]
        quat_w = rs[:, 3:7].clone()
        self.rc_pos_w[ids], self.rc_quat_w[ids] = pos_w, quat_w
        R = self.R
        starts_l = _t(r["ray_starts"]).unsqueeze(0).expand(len(ids), R, 3)
        dirs_l = _t(r["ray_dirs"]).unsqueeze(0).expand(len(ids), R, 3)
        qr = quat_w.unsqueeze(1).expand(len(ids), R, 4)
        if r["attach_yaw_only"]:
            starts_w = mu.quat_apply_yaw(qr.reshape(-1, 4), starts_l.reshape(-1, 3)).view(len(ids), R, 3)
            starts_w = starts_w + pos_w.unsqueeze(1)
            dirs_w = dirs_l
        else:
            starts_w = mu.quat_apply(qr.reshape(-1, 4), starts_l.reshape(-1, 3)).view(len(ids), R, 3) + pos_w.unsqueeze(1)
            dirs_w = mu.quat_apply(qr.reshape(-1, 4), dirs_l.reshape(-1, 3)).view(len(ids), R, 3)
        hits = self.mesh.raycast(starts_w.reshape(-1, 3), dirs_w.reshape(-1, 3), r["max_distance"])
        self.rc_hits[ids] = hits.view(len(ids), R, 3)
        self.rc_ts_last[ids] = self.rc_ts[ids]
        self.rc_outdated[ids] = False

    # ---------------------------------------------------------------------------------------------
    # command term: UniformVelocityCommand (LAB/envs/mdp/commands/velocity_command.py:111-160) under
    # CommandTerm.compute/_resample/reset (LAB/managers/command_manager.py:120-187)
    # ---------------------------------------------------------------------------------------------
    def _resample_command(self, ids, u):
        """u [len(ids), 7] = time_left, x, y, z, heading, is_heading, is_standing draws."""
        c = self.spec["command"]
        lo, hi = c["resampling_time_range"]
        self.cmd_time_left[ids] = mu.scale_uniform(u[:, 0], _t(lo), _t(hi))  # Tensor.uniform_(lo, hi)
        self.cmd_counter[ids] += 1
        rg = c["ranges"]
        for k, key in enumerate(("lin_vel_x", "lin_vel_y", "ang_vel_z")):
            self.cmd_vel_b[ids, k] = mu.scale_uniform(u[:, 1 + k], _t(rg[key][0]), _t(rg[key][1]))
        if c["heading_command"]:
            self.cmd_heading_target[ids] = mu.scale_uniform(u[:, 4], _t(rg["heading"][0]), _t(rg["heading"][1]))
            self.cmd_is_heading[ids] = u[:, 5] <= c["rel_heading_envs"]
        self.cmd_is_standing[ids] = u[:, 6] <= c["rel_standing_envs"]

    def _command_compute(self, d, u_step):
        c = self.spec["command"]
        dt = self.spec["step_dt"]
        max_command_step = c["resampling_time_range"][1] / dt
        self.cmd_metric_xy += torch.norm(self.cmd_vel_b[:, :2] - d["lin_b"][:, :2], dim=-1) / max_command_step
        self.cmd_metric_yaw += torch.abs(self.cmd_vel_b[:, 2] - d["ang_b"][:, 2]) / max_command_step
        self.cmd_time_left -= dt
        ids = (self.cmd_time_left <= 0.0).nonzero().flatten()
        self.last_resample_ids = ids
        if len(ids) > 0:
            self._resample_command(ids, u_step[ids, U_STEP_CMD:U_STEP_CMD + 7])
        if c["heading_command"]:
            hid = self.cmd_is_heading.nonzero(as_tuple=False).flatten()
            err = mu.wrap_to_pi(self.cmd_heading_target[hid] - d["heading_w"][hid])
            self.cmd_vel_b[hid, 2] = torch.clip(c["heading_control_stiffness"] * err, min=c["ranges"]["ang_vel_z"][0],
                                                max=c["ranges"]["ang_vel_z"][1])
        sid = self.cmd_is_standing.nonzero(as_tuple=False).flatten()
        self.cmd_vel_b[sid, :] = 0.0

    # ---------------------------------------------------------------------------------------------
    # ManagerBasedRLEnv._reset_idx (LAB/envs/manager_based_rl_env.py:347-392)
    # ---------------------------------------------------------------------------------------------
    def _reset_idx(self, ids, sim, u):
        s = self.spec
        ur = u[ids]
        log = {}
        # curriculum: terrain_levels_vel (TASKS/.../velocity/mdp/curriculums.py:26-55) + update_env_origins
        # (LAB/terrains/terrain_importer.py:307-322)
        if s.get("curriculum"):
            cur = s["curriculum"]
            pos = sim["root_state_w"][ids, :2]
            dist = torch.norm(pos - self.env_origins[ids, :2], dim=1)
            move_up = dist > cur["terrain_size_x"] / 2
            move_down = dist < torch.norm(self.cmd_vel_b[ids, :2], dim=1) * s["max_episode_length_s"] * 0.5
            move_down *= ~move_up
            self.terrain_levels[ids] += 1 * move_up - 1 * move_down
            max_level = self.terrain_origins.shape[0]
            # torch.randint_like(levels, max_level) made explicit: floor(u * max_level)
            rnd = torch.clamp((ur[:, U_RESET_CURRICULUM] * max_level).to(torch.int64), max=max_level - 1)
            self.terrain_levels[ids] = torch.where(self.terrain_levels[ids] >= max_level, rnd,
                                                   torch.clip(self.terrain_levels[ids], 0))
            self.env_origins[ids] = self.terrain_origins[self.terrain_levels[ids], self.terrain_types[ids]]
        # scene.reset: actuators (actuator_net.py:70-74,142-145)
        for act, st in zip(s["actuators"], self.act_state):
            Jg = len(act["joint_ids"])
            if act["kind"] == "lstm":
                L, H = act["lstm"]["num_layers"], act["lstm"]["hidden"]
                st["h"].view(L, self.N, Jg, H)[:, ids] = 0.0
                st["c"].view(L, self.N, Jg, H)[:, ids] = 0.0
            elif act["kind"] == "mlp":
                st["pos_hist"][ids] = 0.0
                st["vel_hist"][ids] = 0.0
            elif act["kind"] in DELAYED_KINDS:   # new lag = randint(min_delay, max_delay + 1), buffers cleared (actuator_pd.py:311-334)
                k = [a_ for a_ in s["actuators"] if a_["kind"] in DELAYED_KINDS].index(act)
                span = act["max_delay"] - act["min_delay"] + 1
                lag = act["min_delay"] + torch.clamp((ur[:, u_reset_delay(s) + k] * span).to(torch.int64), max=span - 1)
                st["lags"][ids] = lag
                st["ring"].reset(ids)
        for ring in self.obs_hist.values():   # ObservationManager.reset (observation_manager.py:226-230)
            ring.reset(ids)
        # scene.reset: sensors (sensor_base.py:182-195, contact_sensor.py:143-162, ray_caster.py:107-114)
        if s.get("contact"):
            self.cs_ts[ids], self.cs_ts_last[ids], self.cs_outdated[ids] = 0.0, 0.0, True
            self.cs_net_forces_w[ids] = 0.0
            self.cs_hist[ids] = 0.0
            self.cs_cur_air[ids], self.cs_last_air[ids] = 0.0, 0.0
            self.cs_cur_contact[ids], self.cs_last_contact[ids] = 0.0, 0.0
        if s.get("raycaster"):
            r = s["raycaster"]
            self.rc_ts[ids], self.rc_ts_last[ids], self.rc_outdated[ids] = 0.0, 0.0, True
            lo, hi = r["drift_range"]
            self.rc_drift[ids] = mu.scale_uniform(ur[:, U_RESET_DRIFT:U_RESET_DRIFT + 3], _t(lo), _t(hi))
        # reset events in cfg order (LAB/envs/mdp/events.py:823-865, 987-1051)
        ev = s.get("events", {})
        keys = ["x", "y", "z", "roll", "pitch", "yaw"]
        def gated(key, e, env_ids):
            """EventManager.apply(mode="reset") trigger logic (event_manager.py:232-270): the envs of `env_ids` the term
            applies to; min_step_count_between_reset == 0 bypasses the bookkeeping."""
            m = int(e.get("min_step_count", 0) or 0)
            if m == 0:
                return env_ids
            last, once = self._gate.setdefault(key, (torch.zeros(self.N, dtype=torch.int32),
                                                     torch.zeros(self.N, dtype=torch.bool)))
            gstep = self.common_step_counter + 1   # env._sim_step_counter // decimation inside this step
            valid = (gstep - last[env_ids]) >= m
            valid |= (last[env_ids] == 0) & ~once[env_ids]
            vids = env_ids[valid]
            once[vids] = True
            last[vids] = gstep
            return vids

        if ev.get("reset_root_state"):
            e = ev["reset_root_state"]
            vids = gated("root", e, ids)
            if len(vids) > 0:
                uv = u[vids]
                root = self.default_root_state.unsqueeze(0).repeat(len(vids), 1)
                rg = torch.tensor([e["pose_range"].get(k, (0.0, 0.0)) for k in keys], dtype=F32)
                smp = mu.scale_uniform(uv[:, U_RESET_POSE:U_RESET_POSE + 6], rg[:, 0], rg[:, 1])
                positions = root[:, 0:3] + self.env_origins[vids] + smp[:, 0:3]
                dq = mu.quat_from_euler_xyz(smp[:, 3], smp[:, 4], smp[:, 5])
                orientations = mu.quat_mul(root[:, 3:7], dq)
                rg = torch.tensor([e["velocity_range"].get(k, (0.0, 0.0)) for k in keys], dtype=F32)
                smp = mu.scale_uniform(uv[:, U_RESET_VEL:U_RESET_VEL + 6], rg[:, 0], rg[:, 1])
                velocities = root[:, 7:13] + smp
                sim["root_state_w"][vids, 0:3] = positions
                sim["root_state_w"][vids, 3:7] = orientations
                sim["root_state_w"][vids, 7:13] = velocities
        # every reset_joints_by_* term rewrites ALL joints (the reference ignores asset_cfg.joint_ids there)
        for k, e in enumerate(ev.get("reset_joints", [])):
            vids = gated(f"joints{k}", e, ids)
            if len(vids) == 0:
                continue
            uv = u[vids]
            J = self.J
            base = U_RESET_JOINT + 2 * J * k
            jp = self.default_joint_pos.unsqueeze(0).repeat(len(vids), 1)
            jv = self.default_joint_vel.unsqueeze(0).repeat(len(vids), 1)
            up, uvv = uv[:, base:base + J], uv[:, base + J:base + 2 * J]
            (plo, phi), (vlo, vhi) = e["position_range"], e["velocity_range"]
            sp, sv = up * (phi - plo) + plo, uvv * (vhi - vlo) + vlo  # python-float span (math.py:1331)
            if e["mode"] == "scale":
                jp, jv = jp * sp, jv * sv
            else:
                jp, jv = jp + sp, jv + sv
            jp = jp.clamp_(self.soft_joint_pos_limits[:, 0], self.soft_joint_pos_limits[:, 1])
            jv = jv.clamp_(-self.soft_joint_vel_limits, self.soft_joint_vel_limits)
            sim["joint_pos"][vids] = jp  # write_joint_state_to_sim (articulation.py:516,548-550)
            sim["joint_vel"][vids] = jv
            self.prev_joint_vel[vids] = jv
            self.joint_acc[vids] = 0.0
        # manager resets in the order of manager_based_rl_env.py:364-389
        self.prev_action[ids], self.action[ids], self.raw_actions[ids] = 0.0, 0.0, 0.0
        for i, t in enumerate(s["rewards"]):  # reward_manager.py:100-126
            log["Episode_Reward/" + t["name"]] = torch.mean(self.episode_sums[i, ids]) / s["max_episode_length_s"]
            self.episode_sums[i, ids] = 0.0
        if s.get("curriculum"):
            log["Curriculum/terrain_levels"] = torch.mean(self.terrain_levels.float())
        if s.get("command"):  # command_manager.py:120-149
            log["Metrics/base_velocity/error_vel_xy"] = torch.mean(self.cmd_metric_xy[ids]).item()
            log["Metrics/base_velocity/error_vel_yaw"] = torch.mean(self.cmd_metric_yaw[ids]).item()
            self.cmd_metric_xy[ids], self.cmd_metric_yaw[ids] = 0.0, 0.0
            self.cmd_counter[ids] = 0
            self._resample_command(ids, ur[:, U_RESET_CMD:U_RESET_CMD + 7])
        # EventManager.reset (event_manager.py:123-148) only re-draws CLASS-type interval terms: no-op here
        for i, t in enumerate(s["terminations"]):  # termination_manager.py:127-149
            log["Episode_Termination/" + t["name"]] = torch.count_nonzero(self.term_dones[i, ids]).item()
        self.episode_length_buf[ids] = 0
        self.log = log

    # ---------------------------------------------------------------------------------------------
    # ManagerBasedRLEnv.step after the physics loop (manager_based_rl_env.py:198-242)
    # ---------------------------------------------------------------------------------------------
    def post_step(self, sim, u_reset, u_step):
        s = self.spec
        dt = s["step_dt"]
        self.episode_length_buf += 1
        d = self.derived(sim)
        # TerminationManager.compute (termination_manager.py:151-174)
        self.truncated[:] = False
        self.terminated[:] = False
        for i, t in enumerate(s["terminations"]):
            v = self.eval_termination(t, sim, d)
            if t.get("time_out", False):
                self.truncated |= v
            else:
                self.terminated |= v
            self.term_dones[i] = v
        self.reset_buf = self.truncated | self.terminated
        # RewardManager.compute (reward_manager.py:128-157)
        self.reward_buf[:] = 0.0
        for i, t in enumerate(s["rewards"]):
            if t["weight"] == 0.0:
                continue
            value = self.eval_reward(t, sim, d) * t["weight"] * dt
            self.reward_buf += value
            self.episode_sums[i] += value
            self.step_reward[:, i] = value / dt
        ids = self.reset_buf.nonzero(as_tuple=False).squeeze(-1)
        self.last_reset_ids = ids
        if len(ids) > 0:
            self._reset_idx(ids, sim, u_reset)
            self.write_data_to_sim(sim)   # manager_based_rl_env.py:221-223 (scene.write_data_to_sim(); sim.forward())
            d = self.derived(sim)
        if s.get("command"):
            self._command_compute(d, u_step)
        # EventManager.apply(mode="interval") (event_manager.py:204-229) + push_by_setting_velocity (events.py:795-820)
        for k, e in enumerate(s.get("events", {}).get("interval_push", [])):
            if e.get("is_global_time"):   # event_manager.py:213-220: one timer, one draw, every env is pushed
                tl = self.interval_global_time[k]
                tl -= dt
                vid = torch.arange(self.N) if float(tl) < 1e-6 else torch.zeros(0, dtype=torch.long)
                if len(vid) > 0:
                    lo, hi = e["interval_range_s"]
                    tl[:] = u_step[0, U_STEP_INTERVAL + 7 * k] * (hi - lo) + lo
            else:
                tl = self.interval_time_left[k]
                tl -= dt
                vid = (tl < 1e-6).nonzero().flatten()
            if len(vid) > 0:
                base = U_STEP_INTERVAL + 7 * k
                lo, hi = e["interval_range_s"]
                if not e.get("is_global_time"):
                    tl[vid] = u_step[vid, base] * (hi - lo) + lo
                keys = ["x", "y", "z", "roll", "pitch", "yaw"]
                rg = torch.tensor([e["velocity_range"].get(kk, (0.0, 0.0)) for kk in keys], dtype=F32)
                vel = sim["root_state_w"][vid, 7:13]
                vel = vel + mu.scale_uniform(u_step[vid, base + 1:base + 7], rg[:, 0], rg[:, 1])
                sim["root_state_w"][vid, 7:13] = vel

    # ObservationManager.compute_group (LAB/managers/observation_manager.py:260-335) + uniform_noise
    # (LAB/utils/noise/noise_model.py:45-70)
    def compute_obs(self, sim, u_obs):
        o = self.spec["observations"]
        d = self.derived(sim)
        outs, col = [], 0
        for i, t in enumerate(o["terms"]):
            x = self.eval_obs(t, sim, d).clone()
            w = x.shape[1]
            H = max(1, int(t.get("history_length", 0) or 0))
            x = apply_modifiers(x, t.get("modifiers"))   # observation_manager.py:310-312: before the noise
            nz = t.get("noise")
            if nz is not None and o.get("enable_corruption", True):
                u = u_obs[:, col + (H - 1) * w:col + H * w]   # a history term's draw sits at the columns of its newest row
                x = apply_noise(x, u, nz)
            if t.get("clip") is not None:
                x = x.clip_(min=t["clip"][0], max=t["clip"][1])
            if t.get("scale") is not None:
                x = x.mul_(t["scale"])
            if i in self.obs_hist:   # observation_manager.py:319-325: append, then the whole window, oldest first, flattened
                self.obs_hist[i].append(x)
                x = self.obs_hist[i].oldest_first().reshape(self.N, -1)
            outs.append(x)
            col += H * w
        self.obs_buf = torch.cat(outs, dim=-1)
        return self.obs_buf

    def step(self, action, sims, u_reset, u_step, u_obs, sim_prev):
        """One env step.  `sim_prev`: simulator state before the first physics sub-step (the last state of the
        previous step, incl. reset writes); `sims`: the S states after each physics sub-step.  The last one is
        mutated by reset / push events, as the reference mutates PhysX state."""
        self.process_action(action)
        prev = sim_prev
        for sim in sims:
            self.substep_pre(prev)
            self.substep_post(sim)
            prev = sim
        self.post_step(prev, u_reset, u_step)
        self.compute_obs(prev, u_obs)
        self.common_step_counter += 1
        return self.obs_buf, self.reward_buf, self.terminated, self.truncated
