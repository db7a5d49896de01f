"""Recording hook: wraps a running simulator backend (`LiveSimBackend`, envs/manager_based_rl_env.py) and keeps the
PhysX-view tensors it shows the env -- root state, joint state, net contact forces, body velocities, one row per physics
sub-step -- in the slab layout the replay backends (`SyntheticSim`, `HostReplaySim`) and `bench.py` consume:

    root_state_w [K+1, N, 13]   joint_pos / joint_vel [K+1, S, N, J]   net_forces_w [K+1, S, N, B, 3]   body_lin_vel_w [K+1, N, B, 3]

(slab 0 = the state before the first step, as in `generate_slabs`).  A recording of a real Isaac Sim run made this way
replaces the synthetic slabs of the parity tests and the benchmark without touching either: same keys, same shapes.
"""
from __future__ import annotations

import torch


class RecordingSim:
    """Pass-through `LiveSimBackend` that records what the wrapped backend exposes.  `slabs()` after K env steps."""

    def __init__(self, backend, decimation: int, max_steps: int | None = None):
        self.backend, self.S, self.max_steps = backend, int(decimation), max_steps
        self._rows = []          # one dict per physics sub-step, taken right after `physics_step`
        self._first = None       # the state before the first step

    @staticmethod
    def _snap(view):
        return {k: getattr(view, k).detach().clone() for k in ("root_state_w", "joint_pos", "joint_vel", "net_forces_w",
                                                                "body_lin_vel_w") if getattr(view, k, None) is not None}

    def current_view(self):
        v = self.backend.current_view()
        if self._first is None:
            self._first = self._snap(v)
        return v

    def last_view(self):
        return self.backend.last_view()

    def physics_step(self, env):
        if self._first is None:
            self._first = self._snap(self.backend.current_view())
        self.backend.physics_step(env)
        if self.max_steps is None or len(self._rows) < self.max_steps * self.S:
            self._rows.append(self._snap(self.backend.current_view()))

    @property
    def steps_recorded(self) -> int:
        return len(self._rows) // self.S

    def slabs(self) -> dict:
        """The recording as replay slabs (CPU tensors).  Root state and body velocities are kept once per env step (their
        value after the step's last sub-step: what post_step / compute_obs read), joint state and forces per sub-step."""
        K, S = self.steps_recorded, self.S
        if K == 0 or self._first is None:
            raise RuntimeError("nothing recorded yet")
        rows = self._rows[:K * S]

        def per_substep(key):
            first = self._first[key]
            slab0 = torch.stack([first] * S)                                  # the initial state repeated over S
            steps = [torch.stack([rows[k * S + s][key] for s in range(S)]) for k in range(K)]
            return torch.stack([slab0] + steps).cpu()

        def per_step(key):
            return torch.stack([self._first[key]] + [rows[k * S + S - 1][key] for k in range(K)]).cpu()

        out = {"root_state_w": per_step("root_state_w"), "joint_pos": per_substep("joint_pos"),
               "joint_vel": per_substep("joint_vel")}
        if "net_forces_w" in self._first:
            out["net_forces_w"] = per_substep("net_forces_w")
        if "body_lin_vel_w" in self._first:
            out["body_lin_vel_w"] = per_step("body_lin_vel_w")
        return out

    def save(self, path: str):
        torch.save(self.slabs(), path)
