"""Synthetic simulator state: ArticulationData / ContactSensorData-shaped tensors standing in for PhysX.

PhysX is closed and out of scope (north_star); its outputs enter the non-physics step as K pre-generated "slabs"
(one per env step, S physics sub-steps each) drawn on the CPU from a seeded `torch.Generator` -- identical bits for
the CUDA path and the CPU oracle -- and cycled so that consecutive steps do not hit a warm L2.
A recorded Isaac Sim run with the same tensor names can be replayed through `from_recording`.
"""
from __future__ import annotations

import math

import torch


def generate_slabs(spec: dict, K: int, seed: int, foot_body_ids=None) -> dict:
    """CPU tensors: joint_pos/joint_vel [K,S,N,J], net_forces_w [K,S,N,B,3], root_state_w [K,N,13],
    body_lin_vel_w [K,N,B,3]."""
    g = torch.Generator().manual_seed(seed)
    N, J, B, S = spec["N"], spec["J"], spec["B"], spec["decimation"]
    origins = torch.as_tensor(spec["env_origins"], dtype=torch.float32)
    dflt = torch.tensor(spec["default_root_state"], dtype=torch.float32)
    rn = lambda *s: torch.randn(*s, generator=g)  # noqa: E731
    ru = lambda *s: torch.rand(*s, generator=g)  # noqa: E731
    root = torch.zeros(K, N, 13)
    root[..., 0:2] = origins[:, 0:2] + (ru(K, N, 2) - 0.5)
    root[..., 2] = origins[:, 2] + dflt[2] + 0.02 * rn(K, N)
    yaw, roll, pitch = (ru(K, N) * 2 - 1) * math.pi, 0.1 * rn(K, N), 0.1 * rn(K, N)
    cy, sy, cr, sr, cp, sp = (torch.cos(yaw / 2), torch.sin(yaw / 2), torch.cos(roll / 2), torch.sin(roll / 2),
                              torch.cos(pitch / 2), torch.sin(pitch / 2))
    q = torch.stack([cy * cr * cp + sy * sr * sp, cy * sr * cp - sy * cr * sp, cy * cr * sp + sy * sr * cp,
                     sy * cr * cp - cy * sr * sp], dim=-1)
    root[..., 3:7] = q / q.norm(dim=-1, keepdim=True)
    root[..., 7:13] = 0.5 * rn(K, N, 6)
    lim = torch.as_tensor(spec["soft_joint_pos_limits"], dtype=torch.float32)
    jp = torch.tensor(spec["default_joint_pos"], dtype=torch.float32) + 0.2 * rn(K, S, N, J)
    jp = torch.maximum(torch.minimum(jp, lim[:, 1]), lim[:, 0])
    jv = 2.0 * rn(K, S, N, J)
    p_contact = torch.full((B,), 0.02)
    if foot_body_ids is None:
        foot_body_ids = [i for i, n in enumerate(spec.get("body_names", [])) if "foot" in n.lower() or "ankle" in n.lower()]
    p_contact[list(foot_body_ids)] = 0.5
    fz = (ru(K, S, N, B) < p_contact) * (50.0 * rn(K, S, N, B)).abs()
    forces = torch.zeros(K, S, N, B, 3)
    forces[..., 2] = fz
    forces[..., 0:2] = 0.1 * fz.unsqueeze(-1) * rn(K, S, N, B, 2)
    body_vel = 0.3 * rn(K, N, B, 3)
    return {"root_state_w": root, "joint_pos": jp.contiguous(), "joint_vel": jv, "net_forces_w": forces,
            "body_lin_vel_w": body_vel}


def initial_env_state(spec: dict, seed: int) -> dict:
    """Per SURVEY 8d: episode lengths spread over the episode with 1 % at max_len-1, command time_left U(0,10)."""
    g = torch.Generator().manual_seed(seed)
    N = spec["N"]
    ml = spec["max_episode_length"]
    ep = torch.randint(0, ml, (N,), generator=g)
    ep[torch.rand(N, generator=g) < 0.01] = ml - 1
    return {"episode_length_buf": ep.to(torch.int64), "cmd_time_left": torch.rand(N, generator=g) * 10.0,
            "interval_time_left": torch.rand(N, generator=g) * 12.0,
            "interval_global_time_left": 0.03}   # is_global_time terms: one timer, about to fire


def initial_delay_lags(act: dict, init: dict) -> torch.Tensor:
    """Command delay (physics steps) each env of a DelayedPD group starts with: spread over [min_delay, max_delay]."""
    span = act["max_delay"] - act["min_delay"] + 1
    return (init["episode_length_buf"] % span + act["min_delay"]).to(torch.int64)


def init_fused_state(env, init):
    """Loads `initial_env_state` into a FusedEnv (episode counters, command / interval timers)."""
    env.episode_length_buf.copy_(init["episode_length_buf"])
    if env.spec.get("command"):
        env.cmd_time_left.copy_(init["cmd_time_left"])
    for t in env.interval_time_left:
        t.copy_(init["interval_time_left"])
    for gt in getattr(env, "interval_global_time", {}).values():   # is_global_time terms: one timer, both parity slots
        gt.fill_(float(init.get("interval_global_time_left", 0.03)))
    for act, st in zip(env.spec["actuators"], env.act_state):
        if act["kind"] in ("delayed_pd", "remotized_pd"):
            st["lags"].copy_(initial_delay_lags(act, init))


class SyntheticSim:
    """Device-resident slabs + SimViews.  `views(k)` -> (list of S sub-step views for env step k)."""

    def __init__(self, spec: dict, K: int = 32, seed: int = 1234, device="cuda", slabs: dict | None = None):
        from ..envs.fused_env import SimView
        self.spec, self.K = spec, K
        self.S = spec["decimation"]
        cpu = slabs if slabs is not None else generate_slabs(spec, K, seed)
        self.dev = {k: v.to(device).contiguous() for k, v in cpu.items()}
        d = self.dev
        has_contact = spec.get("contact") is not None
        self._views = []
        for k in range(K):
            row = []
            for s in range(self.S):
                row.append(SimView(d["root_state_w"][k], d["joint_pos"][k, s], d["joint_vel"][k, s],
                                   d["net_forces_w"][k, s] if has_contact else None, d["body_lin_vel_w"][k]))
            self._views.append(row)
        self.k = 0

    def views(self, k: int):
        return self._views[k % self.K]

    def bytes(self) -> int:
        return sum(v.numel() * v.element_size() for v in self.dev.values())
