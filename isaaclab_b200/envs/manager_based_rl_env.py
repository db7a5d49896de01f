"""`ManagerBasedRLEnv`: the reference's env interface (LAB/envs/manager_based_rl_env.py:153-242) over the fused kernels.

`step(action) -> (obs_dict, reward, terminated, time_outs, extras)` keeps the reference's order of operations and
its buffer-ownership convention (the same tensor objects are returned every step).  The simulator is a pluggable
backend that exposes PhysX-view-shaped tensors (`SimBackend`): `SyntheticSim` / `HostReplaySim` replay generated or
recorded ArticulationData; a live Isaac Sim would implement `physics_step` with `sim.step()` between the two
sub-step kernels.

Managers are thin views of the fused state with the reference's attribute surface (`reward_manager.compute`,
`termination_manager.terminated`, `command_manager.get_command`, `action_manager.action`, ...).
"""
from __future__ import annotations

import gc

import torch

from .. import _lib
from .compile import compile_env_cfg
from .fused_env import FusedEnv, SimView


class HostReplaySim:
    """Simulator state arriving from the HOST every step (pinned memory -> device staging sets).

    The S physics states of one env step are packed into one pinned slab, so a step's input is ONE H2D copy.  Three
    device staging sets rotate (current / previous / next): while the kernels of env step k read sets k and k-1 on the
    compute stream, step k+1's slab crosses PCIe on a copy stream into the third set (`prefetch_next`, called by the
    env after it has queued the step's kernels).  A recorded source knows its next state in advance; a live
    simulator filling pinned buffers would trigger the copy when its sub-steps are done."""

    _SETS = 3
    _ORDER = ("root_state_w", "joint_pos", "joint_vel", "net_forces_w", "body_lin_vel_w")

    def __init__(self, spec, slabs_cpu: dict, device="cuda", prefetch: bool = True):
        self.spec, self.S = spec, spec["decimation"]
        self.K = slabs_cpu["root_state_w"].shape[0]
        has_contact = spec.get("contact") is not None
        # body velocities cross PCIe only when a term reads them (feet_slide); net forces only with a contact sensor
        needs_body_vel = any(t["op"] == "feet_slide" for t in spec["rewards"])
        self._fields, off = [], 0
        for name in self._ORDER:
            if (name == "body_lin_vel_w" and not needs_body_vel) or (name == "net_forces_w" and not has_contact):
                continue
            v = slabs_cpu[name]
            nbytes = v[0].numel() * v.element_size()
            self._fields.append((name, off, nbytes, tuple(v.shape[1:]), v.dtype))
            off += (nbytes + 255) // 256 * 256
        self.slab_bytes = off
        self.host = torch.empty(self.K, off, dtype=torch.uint8).pin_memory()
        for name, o, nbytes, shape, dtype in self._fields:
            self.host[:, o:o + nbytes].view(dtype).view(self.K, *shape).copy_(slabs_cpu[name])
        self.stage, self.dev, self._views = [], [], []
        for _ in range(self._SETS):
            buf = torch.empty(off, dtype=torch.uint8, device=device)
            st = {name: buf[o:o + nbytes].view(dtype).view(*shape) for name, o, nbytes, shape, dtype in self._fields}
            self.stage.append(buf)
            self.dev.append(st)
            self._views.append([SimView(st["root_state_w"], st["joint_pos"][s], st["joint_vel"][s],
                                        st["net_forces_w"][s] if has_contact else None,
                                        st["body_lin_vel_w"] if needs_body_vel else None)
                                for s in range(self.S)])
        self.h2d_bytes_per_step = sum(f[2] for f in self._fields)
        self.prefetch_enabled = prefetch
        self.copy_stream = torch.cuda.Stream(device=device) if prefetch else None
        self._ready = [torch.cuda.Event() for _ in range(self._SETS)]
        self._free = torch.cuda.Event()
        self._staged = [None] * self._SETS  # which slab index each set holds (or is receiving)
        self.k = -1
        self.slot = 0
        self._copy(0, self.K - 1)  # "previous" state of the first step: the slab before slab 0 in the cycle
        torch.cuda.current_stream().synchronize()

    def _copy(self, slot, k):
        self.stage[slot].copy_(self.host[k], non_blocking=True)
        self._staged[slot] = k

    def mark_free(self):
        """Records, on the compute stream, the point after which the staging set of the NEXT step is free (its last
        reader was the previous step): `prefetch_next(marked=True)` then waits for this point only, so the env can queue
        the current step's kernels first and the copy still overlaps them."""
        self._free.record(torch.cuda.current_stream())

    def prefetch(self, slot, k, marked=False):
        """H2D of slab k into `slot` on the copy stream, after everything already queued on the compute stream (the
        last readers of that set) has finished -- or after the last `mark_free()` when `marked`."""
        if not marked:
            self._free.record(torch.cuda.current_stream())
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self._free)
            self._copy(slot, k)
            self._ready[slot].record(self.copy_stream)

    def advance(self):
        """Views of the next env step's S physics states."""
        self.k = (self.k + 1) % self.K
        self.slot = (self.slot + 1) % self._SETS
        if not self.prefetch_enabled:
            self._copy(self.slot, self.k)
            return self._views[self.slot]
        if self._staged[self.slot] != self.k:  # first call: nothing was prefetched yet
            self.prefetch(self.slot, self.k)
        torch.cuda.current_stream().wait_event(self._ready[self.slot])
        return self._views[self.slot]

    def prefetch_next(self, marked=False):
        """Queue the H2D of the step after the current one (its target set was last read by the previous step)."""
        if self.prefetch_enabled:
            self.prefetch((self.slot + 1) % self._SETS, (self.k + 1) % self.K, marked)

    def last_view(self):
        return self._views[self.slot][-1]


class DeviceReplaySim:
    """Device-resident slabs (SyntheticSim) behind the same `advance()` interface."""

    def __init__(self, synthetic):
        self.syn = synthetic
        self.k = 0
        self.h2d_bytes_per_step = 0

    def advance(self):
        self.k += 1
        return self.syn.views(self.k)

    def last_view(self):
        return self.syn.views(self.k)[-1]


from .manager_views import (ActionManager, CommandManager, CurriculumManager, EventManager, ObservationManager,  # noqa: E402
                            RewardManager, TerminationManager)
from .scene import SceneView  # noqa: E402


class LiveSimBackend:
    """Protocol of a RUNNING simulator behind `ManagerBasedRLEnv.step` (manager_based_rl_env.py:182-196 of the reference):

        current_view() -> SimView   PhysX-view-shaped tensors of the simulator's CURRENT state (root_state_w [N,13],
                                    joint_pos / joint_vel [N,J], net contact forces [N,B,3], body velocities);
        physics_step(env)           one `sim.step()`: consume the efforts / targets the actuator pass left in
                                    `env.fused.joint_effort_target_sim` / `joint_pos_target_sim`, advance, refresh the tensors.

    The env then runs `process_action`, S x (`substep_pre`, `physics_step`, `substep_post`), `post_step`,
    `post_reset_write`, `compute_obs` -- the per-sub-step entry points, since the post-physics states are not known in
    advance.  Replay backends (`advance()` / `last_view()`) use the one-launch fused decimation instead."""

    def current_view(self):
        raise NotImplementedError

    def physics_step(self, env):
        raise NotImplementedError

    def last_view(self):
        return self.current_view()


class ManagerBasedRLEnv:
    def __init__(self, cfg, sim=None, device="cuda", num_envs=None, terrain=None, seed=0, spec=None, use_cuda_graph=False,
                 log_every_step=False):
        """`log_every_step=True`: `extras["log"]` is filled after every step that reset an env, as the reference does
        (one device->host read per step).  Default: the log is accumulated on the device and read once per iteration
        through `episode_log()` (what the runner mirror does) -- same values after the runner's averaging."""
        self.log_every_step = bool(log_every_step)
        self.cfg = cfg
        self.device = torch.device(device)
        self.spec = spec if spec is not None else compile_env_cfg(cfg, num_envs=num_envs, terrain=terrain, seed=seed)
        self.num_envs = self.spec["N"]
        ext_obs = [t for t in self.spec["observations"]["terms"] if t["op"] == "external" and t.get("dim") is None]
        if ext_obs and self.spec.get("raycaster") and "mesh" not in self.spec["raycaster"]:
            from ..sensors.ray_caster import WarpFreeMesh   # one device mesh shared by the probe env and the final one
            r = self.spec["raycaster"]
            r["mesh"] = WarpFreeMesh(r["mesh_vertices"], r["mesh_faces"], cell_size=r.get("cell_size", 0.0))
        for t in ext_obs:
            t["dim"] = 1   # placeholder width for the probe env below
        self.fused = FusedEnv(self.spec, device=device, seed=seed)
        if sim is None:
            from ..sim.synthetic import SyntheticSim
            sim = DeviceReplaySim(SyntheticSim(self.spec, K=8, seed=1234 + seed, device=device))
        self.sim = sim
        if ext_obs:
            # ObservationManager._prepare_terms (observation_manager.py:420-450): the width of a term is the shape of one
            # call.  The probe env provides the buffers the term reads; the final env is built with the real widths.
            self.scene = SceneView(self)
            self._ep_offset = 0
            for t in ext_obs:
                fn = t["func"](t["cfg"], self) if t["is_class"] else t["func"]
                val = fn(self, **t["params"])
                if val.dim() != 2 or val.shape[0] != self.num_envs:
                    raise ValueError(f"observation term '{t['name']}' must return a [num_envs, D] tensor, got {tuple(val.shape)}")
                t["dim"] = int(val.shape[1])
            self.fused.L.b200lab_env_destroy(self.fused.handle)
            self.fused.handle = None
            self.fused = FusedEnv(self.spec, device=device, seed=seed)
        self.step_dt, self.physics_dt = self.spec["step_dt"], self.spec["physics_dt"]
        self.max_episode_length, self.max_episode_length_s = self.spec["max_episode_length"], self.spec["max_episode_length_s"]
        self.reward_manager, self.termination_manager = RewardManager(self), TerminationManager(self)
        self.command_manager, self.action_manager = CommandManager(self), ActionManager(self)
        self.observation_manager = ObservationManager(self)
        self.event_manager, self.curriculum_manager = EventManager(self), CurriculumManager(self)
        self.scene = SceneView(self)
        self._live = hasattr(self.sim, "physics_step")
        self._ep_offset = 0
        self._bind_external_terms()
        self.extras = {}
        self.common_step_counter = 0
        self._group = self.spec["observations"]["group"]
        f = self.fused
        self.reward_buf = f.reward_buf
        self.reset_terminated, self.reset_time_outs, self.reset_buf = f.terminated.view(torch.bool), f.truncated.view(torch.bool), f.reset_buf.view(torch.bool)
        self.obs_buf = {self._group: f.obs_buf}
        self._action_dev = torch.zeros(self.num_envs, f.A, device=device)
        self._prefetch_next = getattr(self.sim, "prefetch_next", None)
        # pipelined result read-back (`step(..., results_out=pinned_slab)`): the small results leave after post_step and
        # each finished range of observation rows leaves while the next range is computed
        self._d2h_stream = None
        self._d2h_events = []
        self._d2h_done = None
        self.result_chunks = 2   # observation ranges of the pipelined read-back (measured: 2 beats 1 and 4, tools/e2e_probe.py)
        self.use_cuda_graph = bool(use_cuda_graph)
        self._graphs, self._graph_warm = {}, False
        if self.use_cuda_graph:
            if self._ext_pre or self._ext_obs or self._live:
                raise ValueError("use_cuda_graph=True needs a replay backend and fused terms only: user (Python) terms "
                                 "and a live simulator run host code between the kernels of a step")
            f.use_device_step_counter()

    # -- user (Python) terms: func(env, **params) -> Tensor, evaluated every step into ext_values (manager_term_cfg.py:29-52)
    def _bind_external_terms(self):
        def bind(t):
            fn = t["func"](t["cfg"], self) if t["is_class"] else t["func"]   # class terms: manager_base.py:393-395
            return (t, fn)
        s = self.spec
        self._ext_pre = [bind(t) for t in list(s["terminations"]) + list(s["rewards"]) if t["op"] == "external"]
        self._ext_obs = [bind(t) for t in s["observations"]["terms"] if t["op"] == "external"]
        self._class_terms = [fn for t, fn in self._ext_pre + self._ext_obs if t["is_class"]]

    def _eval_external_pre(self):
        """Reward / termination terms on the post-physics state, before the step kernel consumes them.  As in the
        reference they see `episode_length_buf` already incremented (manager_based_rl_env.py:200); a reward term that
        reads `termination_manager.terminated` sees the previous step's mask (terminations and rewards share one kernel)."""
        self._ep_offset = 1
        try:
            ext, N = self.fused.ext_values, self.num_envs
            for t, fn in self._ext_pre:
                val = fn(self, **t["params"])
                if val.numel() != N:
                    raise ValueError(f"term '{t['name']}' must return a [num_envs] tensor, got {tuple(val.shape)}")
                ext[t["ext_row"]].copy_(val.reshape(N))
        finally:
            self._ep_offset = 0

    def _eval_external_obs(self):
        ext, N = self.fused.ext_values, self.num_envs
        reset_ids = None
        for t, fn in self._ext_obs:
            val = fn(self, **t["params"])
            mods = t.get("py_modifiers")
            if mods:   # observation_manager.py:310-312 (and :231-233: class modifiers are reset for the envs that reset)
                import inspect
                inst = t.get("_mod_instances")
                if inst is None:   # class modifiers are constructed with the term's output shape (observation_manager.py:456-470)
                    inst = t["_mod_instances"] = [m.func(cfg=m, data_dim=tuple(val.shape), device=self.device)
                                                  if inspect.isclass(m.func) else None for m in mods]
                elif any(i is not None for i in inst):
                    if reset_ids is None:
                        reset_ids = self.reset_buf.nonzero(as_tuple=False).squeeze(-1)
                    if reset_ids.numel() > 0:
                        for i in inst:
                            if i is not None:
                                i.reset(reset_ids)
                for m, i in zip(mods, inst):
                    val = i(val) if i is not None else m.func(val, **m.params)
            if tuple(val.shape) != (N, t["dim"]):
                raise ValueError(f"observation term '{t['name']}' returned {tuple(val.shape)}, expected {(N, t['dim'])}")
            ext[t["ext_row"]:t["ext_row"] + t["dim"]].copy_(val.t())
        if self._class_terms:   # class terms keep per-env state: reset it for the envs this step reset (one host sync)
            ids = self.reset_buf.nonzero(as_tuple=False).squeeze(-1)
            if ids.numel() > 0:
                for term in self._class_terms:
                    term.reset(ids)

    @property
    def episode_length_buf(self):
        buf = self.fused.episode_length_buf
        return buf + self._ep_offset if self._ep_offset else buf

    @episode_length_buf.setter
    def episode_length_buf(self, v):
        self.fused.episode_length_buf.copy_(v)

    @property
    def num_actions(self):
        return self.fused.A

    @property
    def unwrapped(self):
        return self

    def seed(self, seed: int = -1) -> int:
        """Re-seeds the in-kernel Philox streams (ManagerBasedEnv.seed, LAB/envs/manager_based_env.py:239)."""
        import ctypes
        seed = int(seed) if seed is not None and seed >= 0 else int(torch.seed() & 0x7FFFFFFF)
        self.fused._st.seed = seed
        _lib.check(self.fused.L.b200lab_env_update(self.fused.handle, ctypes.byref(self.fused._st)))
        return seed

    def close(self):
        """Releases the native env handle; the object must not be stepped afterwards."""
        f = self.fused
        if getattr(f, "handle", None):
            f.L.b200lab_env_destroy(f.handle)
            f.handle = None

    def _observation_term(self, op, params):
        """An observation identity called directly (`mdp.joint_vel_rel(env, asset_cfg=...)`): its definition on the scene
        views (LAB/envs/mdp/observations.py:33-173,512-531), fresh tensors.  This is the path of terms that carry Python-side
        modifiers; the step's own observations come from `b200lab_compute_obs`."""
        d = self.scene["robot"].data
        cfg = params.get("asset_cfg")
        ids = slice(None) if cfg is None or cfg.joint_ids is None or isinstance(cfg.joint_ids, slice) else list(cfg.joint_ids)
        if op == "base_pos_z":
            return d.root_pos_w[:, 2:3].clone()
        if op == "base_lin_vel":
            return d.root_lin_vel_b.clone()
        if op == "base_ang_vel":
            return d.root_ang_vel_b.clone()
        if op == "projected_gravity":
            return d.projected_gravity_b.clone()
        if op == "joint_pos":
            return d.joint_pos[:, ids].clone()
        if op == "joint_vel":
            return d.joint_vel[:, ids].clone()
        if op == "joint_pos_rel":
            return d.joint_pos[:, ids] - d.default_joint_pos[:, ids]
        if op == "joint_vel_rel":
            return d.joint_vel[:, ids] - d.default_joint_vel[:, ids]
        if op == "last_action":
            return self.action_manager.action.clone()
        if op == "generated_commands":
            return self.command_manager.get_command(params["command_name"]).clone()
        raise NotImplementedError(f"observation term '{op}' is only evaluated inside b200lab_compute_obs (the height scan's "
                                  "ray hits do not exist before it)")

    def evaluate_term(self, fn, *args, **params):
        """Value of a reward / termination term function called directly (`mdp.track_lin_vel_xy_exp(env, ...)`),
        as the term programs evaluated it in the most recent `step` -- the state the reference function would read
        between steps.  The term must be part of this env's cfg with the same params; anything else has no fused
        program to run and raises."""
        kind, op = getattr(fn, "_b200_kind", None), getattr(fn, "_b200_op", None)
        if kind == "obs":
            return self._observation_term(op, params)
        table = {"rew": "rewards", "term": "terminations"}.get(kind)
        if table is None:
            raise NotImplementedError(f"term '{op}' ({kind}) is only evaluated inside the fused step kernels")
        for i, t in enumerate(self.spec[table]):
            if t["op"] != op:
                continue
            if kind == "term":
                return self.fused.term_dones[i].bool()
            if t["weight"] == 0.0:
                raise ValueError(f"reward term '{t['name']}' has weight 0 and is skipped by the reward manager")
            return self.fused.step_reward[:, i] / t["weight"]  # _step_reward = weighted value / dt (reward_manager.py:153)
        raise KeyError(f"term function '{op}' is not part of this environment's {table} cfg")

    def reset(self, seed=None):
        """Initial observations (ManagerBasedEnv.reset, LAB/envs/manager_based_env.py:264): obs of the current state."""
        if self._ext_obs:
            self._eval_external_obs()
        self.fused.compute_obs(self.sim.last_view())
        return self.obs_buf, self.extras

    def _setup_d2h(self):
        self._d2h_stream = torch.cuda.Stream(device=self.device)
        self._d2h_events = [torch.cuda.Event() for _ in range(self.result_chunks + 2)]
        N, C = self.num_envs, max(1, min(self.result_chunks, self.num_envs))
        edges = [N * i // C for i in range(C + 1)]
        self._d2h_ranges = [(edges[i], edges[i + 1]) for i in range(C) if edges[i + 1] > edges[i]]

    results_mode = "all"   # "all": the whole result slab is read back; "metrics": everything but the observation rows

    def wait_results(self):
        """Blocks until the `results_out` slab of the last `step` is complete on the host."""
        if self._d2h_done is not None:
            self._d2h_done.synchronize()

    def _step_body(self, action, views, prev, results_out, pipelined):
        """The device work of one step on the current stream: stage a host action, the fused kernels, and -- when a
        pinned `results_out` slab is given -- the read-back.  `pipelined`: reward / masks leave right after post_step and
        each finished range of observation rows leaves on a copy stream while the next range is computed (used inside
        a captured graph, where the extra enqueue calls cost nothing at replay); otherwise one copy at the end."""
        f = self.fused
        if results_out is None or not pipelined:
            f.step(action, views, prev, before_post_step=self._eval_external_pre if self._ext_pre else None,
                   after_post_step=self._eval_external_obs if (self._ext_obs or self._class_terms) else None)
            if results_out is not None:
                lo = f.result_layout["reward"][0] if self.results_mode == "metrics" else 0
                results_out[lo:].copy_(f.result[lo:], non_blocking=True)
            return
        if self._d2h_stream is None:
            self._setup_d2h()
        cur, cs, it = torch.cuda.current_stream(), self._d2h_stream, iter(self._d2h_events)
        row = f.obs_buf.shape[1] * 4
        o_rew = f.result_layout["reward"][0]

        def push(lo, hi):
            e = next(it)
            e.record(cur)
            cs.wait_event(e)
            with torch.cuda.stream(cs):
                results_out[lo:hi].copy_(f.result[lo:hi], non_blocking=True)

        if self.results_mode == "metrics":
            # reward / terminated / time-outs / reset mask only: the observations stay on the device for a policy that
            # runs there (what the reference's runner does); the slab's observation part is left untouched
            f.step(action, views, prev, after_post_step=lambda: push(o_rew, f.result.numel()))
        else:
            f.step(action, views, prev, obs_ranges=self._d2h_ranges,
                   after_post_step=lambda: push(o_rew, f.result.numel()),
                   after_range=lambda b, e: push(b * row, e * row))
        e = next(it)
        e.record(cs)
        cur.wait_event(e)   # join: everything is ordered on the current stream again

    def step(self, action: torch.Tensor, results_out: torch.Tensor | None = None):
        """`results_out`: optional pinned uint8 host slab of `fused.result.numel()` bytes (layout `fused.result_layout`,
        typed views from `result_views`) that receives observations, reward, terminated, time-outs and the reset mask;
        call `wait_results()` before reading it.  With `use_cuda_graph=True` the whole step -- action staging, kernels,
        pipelined read-back -- is captured once per (action buffer, result slab, simulator staging set) and replayed
        with one launch; the first step runs eagerly."""
        if action.device != self.device:  # host action: staged through a fixed device buffer
            self._action_dev.copy_(action, non_blocking=True)
            action = self._action_dev
        if self._live:
            return self._step_live(action, results_out)
        prev = self.sim.last_view()
        views = self.sim.advance()
        # The staging set of the NEXT step was last read by the previous step, which precedes everything queued below on
        # this stream: its H2D copy only has to wait for this point (after the small action copy -- H2D copies share one
        # engine and run in issue order) and crosses PCIe while this step's kernels run.  The copy itself is queued after
        # the kernels so that the GPU is not idle while the host enqueues it.
        mark = getattr(self.sim, "mark_free", None) if self._prefetch_next is not None else None
        if mark is not None:
            mark()
        graph = None
        if self.use_cuda_graph:
            key = (action.data_ptr(), -1 if results_out is None else results_out.data_ptr(),
                   tuple(id(v) for v in views), id(prev), self.results_mode)
            graph = self._graphs.get(key)
            if graph is None and self._graph_warm:
                if len(self._graphs) >= 64:
                    self._graphs.clear()
                graph = torch.cuda.CUDAGraph()
                # a collection that frees another env / mesh (cudaFree) while the stream is capturing would invalidate
                # the capture
                gc_was_enabled = gc.isenabled()
                gc.collect()
                gc.disable()
                host_counter = self.fused.step_counter   # the captured body advances it once; the replay below counts
                try:
                    # thread_local: other threads of the process (e.g. the NCCL watchdog polling its events) must not
                    # invalidate the capture
                    with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                        self._step_body(action, views, prev, results_out, pipelined=True)
                finally:
                    self.fused.step_counter = host_counter
                    if gc_was_enabled:
                        gc.enable()
                self._graphs[key] = graph
        if graph is not None:
            graph.replay()
            self.fused.step_counter += 1
        else:
            self._step_body(action, views, prev, results_out, pipelined=False)
            self._graph_warm = True
        if results_out is not None:
            if self._d2h_done is None:
                self._d2h_done = torch.cuda.Event()
            self._d2h_done.record(torch.cuda.current_stream())
        if self._prefetch_next is not None:
            if mark is not None:
                self._prefetch_next(marked=True)
            else:
                self._prefetch_next()
        self.common_step_counter += 1
        if self.log_every_step:
            log = self.fused.last_step_log()
            if log:
                self.extras["log"] = log
            else:
                self.extras.pop("log", None)
        return self.obs_buf, self.reward_buf, self.reset_terminated, self.reset_time_outs, self.extras

    def _step_live(self, action, results_out):
        """One env step around a RUNNING simulator (`LiveSimBackend`): the order of manager_based_rl_env.py:173-236."""
        f, sim = self.fused, self.sim
        f.process_action(action)
        for _ in range(self.spec["decimation"]):
            f.substep_pre(sim.current_view())     # apply_action + actuator models on the pre-physics state
            sim.physics_step(self)                # sim.step()
            f.substep_post(sim.current_view())    # scene.update(dt): joint_acc, contact sensor, sensor clocks
        view = sim.current_view()
        if self._ext_pre:
            self._eval_external_pre()
        f.post_step(view)
        if self._ext_obs or self._class_terms:
            self._eval_external_obs()
        f.post_reset_write(view)
        f.log_fold()
        f.compute_obs(view)
        f.step_counter += 1
        if results_out is not None:
            results_out.copy_(f.result, non_blocking=True)
            if self._d2h_done is None:
                self._d2h_done = torch.cuda.Event()
            self._d2h_done.record(torch.cuda.current_stream())
        self.common_step_counter += 1
        if self.log_every_step:
            log = f.last_step_log()
            if log:
                self.extras["log"] = log
            else:
                self.extras.pop("log", None)
        return self.obs_buf, self.reward_buf, self.reset_terminated, self.reset_time_outs, self.extras

    def result_views(self, host_slab: torch.Tensor):
        """Typed views of a result slab filled by `step(..., results_out=host_slab)` or `read_results`."""
        lay, N = self.fused.result_layout, self.num_envs
        return {"obs": host_slab[lay["obs"][0]:lay["obs"][1]].view(torch.float32).view(N, -1),
                "reward": host_slab[lay["reward"][0]:lay["reward"][1]].view(torch.float32),
                **{k: host_slab[lay[k][0]:lay[k][1]] for k in ("terminated", "truncated", "reset_buf")}}

    def read_results(self, host_slab: torch.Tensor):
        """ONE device->host copy (async on the current stream) of everything `step` returned -- observations, reward,
        terminated, time-outs, reset mask -- into a pinned uint8 slab of `fused.result.numel()` bytes; returns typed host
        views of it (valid after the stream is synchronised)."""
        f = self.fused
        host_slab.copy_(f.result, non_blocking=True)
        lay = f.result_layout
        N = self.num_envs
        return {"obs": host_slab[lay["obs"][0]:lay["obs"][1]].view(torch.float32).view(N, -1),
                "reward": host_slab[lay["reward"][0]:lay["reward"][1]].view(torch.float32),
                **{k: host_slab[lay[k][0]:lay[k][1]] for k in ("terminated", "truncated", "reset_buf")}}

    def episode_log(self):
        """Mean of the per-step extras["log"] dicts (manager_based_rl_env.py:365-389) since the last call: the value the
        reference runner logs for an iteration, read from the device ONCE instead of `.item()` syncs at every reset."""
        return self.fused.mean_log()
