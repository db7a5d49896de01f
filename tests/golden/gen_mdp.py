"""Golden fixtures of the non-physics env step, produced by the UNMODIFIED reference.

Run in the authoring container (needs /root/reference):  python -m tests.golden.gen_mdp
Writes tests/golden/mdp_<task>.pt for the five BASELINE task ids.

What runs here is the reference's own code: `ManagerBasedRLEnv.step` / `_reset_idx`, the real Action/Observation/
Reward/Termination/Command/Event/Curriculum managers built from the reference's own task cfg classes, the real
`Articulation.write_data_to_sim` / `_apply_actuator_model` with the real actuator classes (incl. TorchScript nets),
`ArticulationData`, `ContactSensor`, `RayCaster`, `TerrainImporter.update_env_origins`, `InteractiveScene` dispatch.
Only the closed parts are replaced:
  * PhysX tensor views  -> `ReplayArticulationView` / `ReplayContactView` serving pre-generated simulator state
    (isaaclab_b200.sim.synthetic.generate_slabs -- the same tensors the CUDA path and the oracle consume);
  * USD/Kit discovery (`_initialize_impl`) -> buffers allocated here with the same shapes, zero-filled;
  * warp `raycast_mesh` (warp-lang is not in the image) -> oracle.raycast brute force.  The height-scan columns of the
    fixtures therefore pin everything AROUND the ray query (sensor pose, yaw-only rays, drift, clip) but not the query.
Random draws are RECORDED, not replaced: torch.rand / rand_like / Tensor.uniform_ / randint_like are wrapped so that
each draw is logged with the reference function that made it and that function's `env_ids`; the log is then laid out
in the (u_reset, u_step, u_obs) slot format of include/b200lab.h so the same uniforms can be replayed by the oracle and
the CUDA path.  `Tensor.uniform_(a, b)` is evaluated as u * (b - a) + a in fp32, ATen's own formula.
"""
from __future__ import annotations

import copy
import os
import sys
import tempfile

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from tests.golden import ref_harness  # noqa: E402

ref_harness.install()

import isaaclab.utils.math as ref_math  # noqa: E402
import omni.physics.tensors.impl.api as physx  # noqa: E402  (stub module; classes used for isinstance checks only)
from isaaclab.assets.articulation.articulation import Articulation  # noqa: E402
from isaaclab.assets.articulation.articulation_data import ArticulationData  # noqa: E402
from isaaclab.envs.manager_based_rl_env import ManagerBasedRLEnv  # noqa: E402
from isaaclab.managers import (ActionManager, CommandManager, CurriculumManager, EventManager,  # noqa: E402
                               ObservationManager, RecorderManager, RewardManager, TerminationManager)
from isaaclab.scene.interactive_scene import InteractiveScene  # noqa: E402
from isaaclab.sensors.contact_sensor.contact_sensor import ContactSensor  # noqa: E402
from isaaclab.sensors.contact_sensor.contact_sensor_data import ContactSensorData  # noqa: E402
from isaaclab.sensors.ray_caster import ray_caster as ref_ray_caster  # noqa: E402
from isaaclab.sensors.ray_caster.ray_caster_data import RayCasterData  # noqa: E402
from isaaclab.terrains.terrain_importer import TerrainImporter  # noqa: E402

from oracle import raycast as oracle_raycast  # noqa: E402
from oracle import mdp_env as omdp  # noqa: E402
from tests import mdp_common  # noqa: E402
from isaaclab_b200.sim.synthetic import generate_slabs, initial_env_state  # noqa: E402

REF = ref_harness.REFERENCE_ROOT


# ------------------------------------------------------------------------------------------------------------------
# replayed simulator state behind the PhysX tensor-view surface
# ------------------------------------------------------------------------------------------------------------------
class Replay:
    """Cursor over the slabs: state after physics sub-step s of env step k.  Event writes mutate the slab in place,
    exactly as the reference mutates PhysX state."""

    def __init__(self, slabs, wrap=False):
        self.slabs = slabs
        self.S = slabs["joint_pos"].shape[1]
        self.K = slabs["joint_pos"].shape[0]
        self.wrap = wrap                 # timing runs cycle through the slabs; fixture generation never wraps
        self.k, self.s = 0, self.S - 1   # "before the first sub-step" = last state of slab 0

    def advance(self):
        self.s += 1
        if self.s == self.S:
            self.k, self.s = self.k + 1, 0
            if self.wrap:
                self.k %= self.K

    @property
    def root(self):
        return self.slabs["root_state_w"][self.k]

    @property
    def joint_pos(self):
        return self.slabs["joint_pos"][self.k, self.s]

    @property
    def joint_vel(self):
        return self.slabs["joint_vel"][self.k, self.s]

    @property
    def forces(self):
        return self.slabs["net_forces_w"][self.k, self.s]

    @property
    def body_lin_vel(self):
        return self.slabs["body_lin_vel_w"][self.k]


class _Meta:
    def __init__(self, dof_names, link_names):
        self.dof_names, self.link_names, self.fixed_base = dof_names, link_names, False
        self.dof_count, self.link_count = len(dof_names), len(link_names)


def _new(cls):
    return object.__new__(cls)


class ReplayArticulationView(physx.ArticulationView):
    """The subset of omni.physics.tensors ArticulationView the hot path touches.  Quaternions are xyzw on this side of
    the boundary, as in PhysX."""

    def setup(self, replay, N, joint_names, body_names, limits):
        self.r, self.count, self.max_dofs, self.max_links, self.max_fixed_tendons = replay, N, len(joint_names), len(body_names), 0
        self.shared_metatype = _Meta(list(joint_names), list(body_names))
        self._limits = limits.unsqueeze(0).repeat(N, 1, 1).clone()
        self.written = {}
        return self

    # -- state getters
    def get_root_transforms(self):
        p = self.r.root[:, 0:7].clone()
        p[:, 3:7] = ref_math.convert_quat(p[:, 3:7], to="xyzw")
        return p

    def get_root_velocities(self):
        return self.r.root[:, 7:13].clone()

    def get_dof_positions(self):
        return self.r.joint_pos.clone()

    def get_dof_velocities(self):
        return self.r.joint_vel.clone()

    def get_link_transforms(self):
        t = torch.zeros(self.count, self.max_links, 7)
        t[..., 6] = 1.0
        return t

    def get_link_velocities(self):
        v = torch.zeros(self.count, self.max_links, 6)
        v[..., 0:3] = self.r.body_lin_vel
        return v

    def get_link_accelerations(self):
        return torch.zeros(self.count, self.max_links, 6)

    def get_coms(self):
        t = torch.zeros(self.count, self.max_links, 7)
        t[..., 6] = 1.0
        return t

    # -- property getters (USD defaults; none of them enters the arithmetic of the path for the five tasks)
    def get_dof_limits(self):
        return self._limits.clone()

    def _zeros(self):
        return torch.zeros(self.count, self.max_dofs)

    get_dof_stiffnesses = get_dof_dampings = get_dof_armatures = get_dof_friction_coefficients = _zeros

    def get_dof_max_velocities(self):
        return torch.full((self.count, self.max_dofs), 1.0e9)

    def get_dof_max_forces(self):
        return torch.full((self.count, self.max_dofs), 1.0e9)

    def get_masses(self):
        return torch.ones(self.count, self.max_links)

    def get_inertias(self):
        return torch.ones(self.count, self.max_links, 9)

    # -- state setters: write into the replayed state
    def set_root_transforms(self, data, indices):
        d = data[indices].clone()
        d[:, 3:7] = ref_math.convert_quat(d[:, 3:7], to="wxyz")
        self.r.root[indices, 0:7] = d

    def set_root_velocities(self, data, indices):
        self.r.root[indices, 7:13] = data[indices]

    def set_dof_positions(self, data, indices):
        self.r.joint_pos[indices] = data[indices]

    def set_dof_velocities(self, data, indices):
        self.r.joint_vel[indices] = data[indices]

    # -- command / property setters: remembered, not simulated
    def _remember(name):  # noqa: N805
        def f(self, data, indices=None, *a, **k):
            self.written[name] = data.clone() if isinstance(data, torch.Tensor) else data
        return f

    set_dof_actuation_forces = _remember("effort")
    set_dof_position_targets = _remember("pos_target")
    set_dof_velocity_targets = _remember("vel_target")
    set_dof_stiffnesses = _remember("stiffness")
    set_dof_dampings = _remember("damping")
    set_dof_max_forces = _remember("max_forces")
    set_dof_max_velocities = _remember("max_vel")
    set_dof_armatures = _remember("armature")
    set_dof_friction_coefficients = _remember("friction")
    set_dof_limits = _remember("limits")

    def apply_forces_and_torques_at_position(self, **k):
        self.written["wrench"] = True


class ReplayContactView:
    def setup(self, replay, N, B, body_names):
        self.r, self.count, self.filter_count = replay, N * B, 0
        self.prim_paths = [f"/World/envs/env_0/Robot/{n}" for n in body_names]
        return self

    def get_net_contact_forces(self, dt):
        return self.r.forces.reshape(-1, 3).clone()


class ReplaySim:
    """SimulationContext surface used by ManagerBasedRLEnv.step and ManagerBase."""
    device = "cpu"

    def __init__(self, replay, physics_dt):
        self.r, self._dt = replay, physics_dt

    def is_playing(self):
        return True

    def has_gui(self):
        return False

    def has_rtx_sensors(self):
        return False

    def get_physics_dt(self):
        return self._dt

    def step(self, render=False):
        self.r.advance()

    def forward(self):
        pass

    def render(self):
        pass


def _strict_getattr(self, name):
    raise AttributeError(f"{type(self).__name__} (PhysX stand-in) has no '{name}': the reference touched a view "
                         "method the replay does not provide")


ReplayArticulationView.__getattr__ = _strict_getattr


# ------------------------------------------------------------------------------------------------------------------
# recording of the reference's random draws
# ------------------------------------------------------------------------------------------------------------------
class DrawLog:
    """Wraps the torch RNG entry points the path uses; every draw made from reference code is logged as
    (function name, that frame's env_ids, uniform tensor, names of enclosing reference functions)."""

    _SKIP_FILES = ("utils/math.py",)   # sample_uniform & co: attribute the draw to their caller

    def __init__(self, seed):
        self.gen = torch.Generator().manual_seed(seed)
        self.draws = []
        self.enabled = False
        self._orig = {}

    def _site(self):
        f = sys._getframe(2)
        first, chain, event = None, [], None
        while f is not None:
            fn = f.f_code.co_filename
            if fn.startswith(REF):
                if first is None and not fn.endswith(self._SKIP_FILES):
                    first = f
                chain.append(f.f_code.co_name)
                if f.f_code.co_name == "apply" and fn.endswith("event_manager.py") and "index" in f.f_locals:
                    event = (int(f.f_locals["index"]), bool(f.f_locals["term_cfg"].is_global_time))   # (term index, global timer)
            f = f.f_back
        if first is None:
            return None
        loc = first.f_locals
        ids = loc.get("valid_env_ids", loc.get("env_ids", None))
        if isinstance(ids, torch.Tensor):
            ids = ids.clone()
        return first.f_code.co_name, ids, tuple(chain), event

    def _log(self, u, extra=None):
        if not self.enabled:
            return
        site = self._site()
        if site is not None:
            self.draws.append(dict(func=site[0], env_ids=site[1], chain=site[2], event=site[3], u=u.clone(), extra=extra))

    def install(self):
        log = self
        o_rand, o_rand_like, o_uniform, o_randint_like = torch.rand, torch.rand_like, torch.Tensor.uniform_, torch.randint_like
        self._orig = dict(rand=o_rand, rand_like=o_rand_like, uniform_=o_uniform, randint_like=o_randint_like)

        def rand(*size, **kw):
            kw.pop("device", None)
            if "generator" in kw:
                return o_rand(*size, **kw)
            u = o_rand(*size, generator=log.gen, **kw)
            log._log(u)
            return u

        def rand_like(x, **kw):
            u = o_rand(x.shape, generator=log.gen, dtype=x.dtype)
            log._log(u)
            return u

        def uniform_(self, a=0.0, b=1.0, **kw):
            if "generator" in kw:
                return o_uniform(self, a, b, **kw)
            u = o_rand(self.shape, generator=log.gen, dtype=torch.float32)
            log._log(u)
            lo, hi = torch.tensor(a, dtype=torch.float32), torch.tensor(b, dtype=torch.float32)
            return self.copy_(u * (hi - lo) + lo)

        def randint_like(x, high, **kw):
            r = torch.randint(0, high, x.shape, generator=log.gen, dtype=x.dtype)
            log._log((r.to(torch.float32) + 0.5) / high, extra=("randint", high))
            return r

        o_randint = torch.randint
        self._orig["randint"] = o_randint

        def randint(*a, **kw):   # torch.randint(low=, high=, size=, ...) as DelayedPDActuator.reset calls it
            if "generator" in kw or not log.enabled or a:
                return o_randint(*a, **kw)
            low, high, size = kw.get("low", 0), kw["high"], kw["size"]
            r = o_randint(low, high, size, generator=log.gen, dtype=kw.get("dtype", torch.int64))
            log._log((r.to(torch.float32) - low + 0.5) / (high - low), extra=("randint", low, high))
            return r

        torch.rand, torch.rand_like, torch.Tensor.uniform_, torch.randint_like = rand, rand_like, uniform_, randint_like
        torch.randint = randint

    def uninstall(self):
        o = self._orig
        torch.rand, torch.rand_like, torch.Tensor.uniform_, torch.randint_like = (o["rand"], o["rand_like"], o["uniform_"],
                                                                                 o["randint_like"])
        torch.randint = o["randint"]

    def take(self):
        d, self.draws = self.draws, []
        return d


def layout_uniforms(draws, spec, N, filler):
    """Recorded draws of ONE env step -> (u_reset, u_step, u_obs) in the slot layout of include/b200lab.h.
    Slots the reference did not draw this step keep `filler` values (they must not be read)."""
    J = spec["J"]
    n_r, n_s, n_o = omdp.n_u_reset(spec), omdp.n_u_step(spec), omdp.obs_dim(spec)
    u_r, u_s, u_o = filler(N, n_r), filler(N, n_s), filler(N, n_o)
    seen = {}
    obs_col = 0
    noisy_cols = []   # (start, width) of obs terms that carry noise, in term order
    c = 0
    for t in spec["observations"]["terms"]:
        w, wt = omdp.term_raw_dim(spec, t), omdp.term_dim(spec, t)
        if t.get("noise") is not None and spec["observations"].get("enable_corruption", True):
            noisy_cols.append((c + wt - w, w))   # a history term's draw sits at the columns of its newest row
        c += wt
    unused = []
    for d in draws:
        f, ids, chain, u = d["func"], d["env_ids"], d["chain"], d["u"]
        k = seen.get((f, "reset" if "_reset_idx" in chain else "step"), 0)
        seen[(f, "reset" if "_reset_idx" in chain else "step")] = k + 1
        in_reset = "_reset_idx" in chain
        if isinstance(ids, slice) or ids is None:
            ids = torch.arange(N)
        if f == "update_env_origins":
            u_r[ids, omdp.U_RESET_CURRICULUM] = u
        elif f == "reset" and d["extra"] is not None and d["extra"][0] == "randint":
            # DelayedPDActuator.reset (actuator_pd.py:320-326): one lag per env, groups in articulation order
            n_delayed = sum(1 for a_ in spec["actuators"] if a_["kind"] in ("delayed_pd", "remotized_pd"))
            kk = seen.get("delay_total", 0)
            seen["delay_total"] = kk + 1
            u_r[ids, omdp.u_reset_delay(spec) + kk % n_delayed] = u
        elif f == "reset":   # RayCaster.reset (the only sensor reset that draws)
            u_r[ids, omdp.U_RESET_DRIFT:omdp.U_RESET_DRIFT + 3] = u
        elif f == "reset_root_state_uniform":
            base = omdp.U_RESET_POSE if k == 0 else omdp.U_RESET_VEL
            u_r[ids, base:base + 6] = u
        elif f in ("reset_joints_by_scale", "reset_joints_by_offset"):
            kk = seen.get("reset_joints_total", 0)
            seen["reset_joints_total"] = kk + 1
            ev, half = divmod(kk, 2)
            base = omdp.U_RESET_JOINT + 2 * J * ev + J * half
            u_r[ids, base:base + J] = u
        elif f == "_resample":
            (u_r if in_reset else u_s)[ids, omdp.U_RESET_CMD if in_reset else omdp.U_STEP_CMD] = u
        elif f == "_resample_command":
            heading = spec["command"]["heading_command"]
            order = [1, 2, 3] + ([4, 5] if heading else []) + [6]
            slot = order[k % len(order)]
            (u_r if in_reset else u_s)[ids, (omdp.U_RESET_CMD if in_reset else omdp.U_STEP_CMD) + slot] = u
        elif f == "apply":   # EventManager.apply(mode="interval"): new time_left of interval term `index` this step
            index, is_global = d["event"]
            base = omdp.U_STEP_INTERVAL + 7 * index
            if is_global:    # torch.rand(1), one timer for all envs (event_manager.py:213-217): env 0's slot
                u_s[0, base] = u[0]
            else:
                u_s[ids, base] = u
        elif f == "push_by_setting_velocity":
            index, is_global = d["event"]
            base = omdp.U_STEP_INTERVAL + 7 * index
            u_s[torch.arange(N) if is_global else ids, base + 1:base + 7] = u
        elif f == "uniform_noise":
            s0, w = noisy_cols[obs_col]
            assert u.shape == (N, w), (u.shape, w)
            u_o[:, s0:s0 + w] = u
            obs_col += 1
        else:
            unused.append(f)
    return u_r, u_s, u_o, unused


# ------------------------------------------------------------------------------------------------------------------
# reference objects, constructed without Kit
# ------------------------------------------------------------------------------------------------------------------
def script_lstm(net, path):
    """TorchScript file with the structure ActuatorNetLSTM expects (`.lstm` + Linear head), holding `net`'s weights."""
    class SeaNet(torch.nn.Module):
        def __init__(self, L, H):
            super().__init__()
            self.lstm = torch.nn.LSTM(2, H, L, batch_first=True)
            self.fc = torch.nn.Linear(H, 1)

        def forward(self, x, hc: tuple[torch.Tensor, torch.Tensor]):
            y, hc2 = self.lstm(x, hc)
            return self.fc(y[:, -1]), hc2

    m = SeaNet(net["num_layers"], net["hidden"])
    sd = {f"lstm.{k}": v for k, v in net.items() if k.startswith(("weight_", "bias_"))}
    sd["fc.weight"], sd["fc.bias"] = net["head_weight"], net["head_bias"]
    m.load_state_dict(sd)
    torch.jit.script(m.eval()).save(path)


def script_mlp(net, path):
    acts = {"softsign": torch.nn.Softsign, "elu": torch.nn.ELU, "tanh": torch.nn.Tanh, "relu": torch.nn.ReLU}
    layers = []
    for i, (w, b) in enumerate(net["layers"]):
        lin = torch.nn.Linear(w.shape[1], w.shape[0])
        lin.weight.data.copy_(w)
        lin.bias.data.copy_(b)
        layers.append(lin)
        if i + 1 < len(net["layers"]):
            layers.append(acts[net["activation"]]())
    torch.jit.script(torch.nn.Sequential(*layers).eval()).save(path)


class _PhysicsSimView:
    def get_gravity(self):
        return (0.0, 0.0, -9.81)   # SimulationCfg.gravity default (LAB/sim/simulation_cfg.py)

    def update_articulations_kinematic(self):
        pass


class _SimulationManager:
    @staticmethod
    def get_physics_sim_view():
        return _PhysicsSimView()


def _quiet_handles(obj):
    obj._initialize_handle = obj._invalidate_initialize_handle = obj._debug_vis_handle = None


def build_articulation(ref_cfg, spec, replay, tmpdir):
    import isaaclab.assets.articulation.articulation_data as ad
    ad.SimulationManager = _SimulationManager
    N = spec["N"]
    limits = torch.as_tensor(spec["joint_pos_limits"], dtype=torch.float32)
    view = _new(ReplayArticulationView).setup(replay, N, spec["joint_names"], spec["body_names"], limits)
    cfg = ref_cfg.copy()
    for (name, acfg), aspec in zip(cfg.actuators.items(), spec["actuators"]):
        if aspec["kind"] == "lstm":
            acfg.network_file = os.path.join(tmpdir, f"{name}_lstm.pt")
            script_lstm(aspec["lstm"], acfg.network_file)
        elif aspec["kind"] == "mlp":
            acfg.network_file = os.path.join(tmpdir, f"{name}_mlp.pt")
            script_mlp(aspec["mlp"], acfg.network_file)
    a = _new(Articulation)
    _quiet_handles(a)
    a.cfg, a._device, a._is_initialized = cfg, "cpu", True
    a._root_physx_view = view
    a._data = ArticulationData(view, "cpu")
    a._data._previous_joint_vel.zero_()   # joint_acc history starts from rest on both sides
    a._create_buffers()
    a._process_cfg()
    a._process_actuators_cfg()
    a._fixed_tendon_names = []
    a._data.fixed_tendon_names = []
    a.update(0.0)
    return a


def build_contact_sensor(cfg, spec, replay):
    N, B = spec["N"], spec["B"]
    s = _new(ContactSensor)
    _quiet_handles(s)
    s.cfg = cfg.copy()
    s._is_initialized, s._is_visualizing, s._device, s._num_envs, s._num_bodies = True, False, "cpu", N, B
    s._sim_physics_dt = spec["physics_dt"]
    s._is_outdated = torch.ones(N, dtype=torch.bool)
    s._timestamp = torch.zeros(N)
    s._timestamp_last_update = torch.zeros(N)
    s._contact_physx_view = ReplayContactView().setup(replay, N, B, spec["body_names"])
    s._body_physx_view = s._contact_physx_view
    d = s._data = ContactSensorData()
    H = s.cfg.history_length
    d.net_forces_w = torch.zeros(N, B, 3)
    d.net_forces_w_history = torch.zeros(N, H, B, 3) if H > 0 else d.net_forces_w.unsqueeze(1)
    if s.cfg.track_air_time:
        d.last_air_time, d.current_air_time = torch.zeros(N, B), torch.zeros(N, B)
        d.last_contact_time, d.current_contact_time = torch.zeros(N, B), torch.zeros(N, B)
    return s


class _OracleMeshHandle:
    def __init__(self, mesh):
        self.mesh = mesh


def _raycast_mesh_stand_in(ray_starts, ray_directions, mesh, max_dist=1e6, **kw):
    """Replaces LAB/utils/warp/ops.py:raycast_mesh (warp-lang absent): same contract -- hit positions, inf on miss."""
    shape = ray_starts.shape
    hits = mesh.mesh.raycast(ray_starts.reshape(-1, 3).contiguous(), ray_directions.reshape(-1, 3).contiguous(), max_dist)
    return hits.view(shape), None, None, None


def build_ray_caster(cfg, spec, art_view):
    N = spec["N"]
    s = _new(ref_ray_caster.RayCaster)
    _quiet_handles(s)
    s.cfg = cfg.copy()
    s._is_initialized, s._is_visualizing, s._device, s._num_envs = True, False, "cpu", N
    s._is_outdated = torch.ones(N, dtype=torch.bool)
    s._timestamp = torch.zeros(N)
    s._timestamp_last_update = torch.zeros(N)
    s._data = RayCasterData()
    s._view = art_view
    r = spec["raycaster"]
    mesh = oracle_raycast.Mesh(r["mesh_vertices"], r["mesh_faces"])
    if mesh.F.shape[0] > 200_000:   # full-size terrain (timing runs): exact xy pre-filter, same hits as brute force
        mesh.enable_grid(0.25)
    s.meshes = {s.cfg.mesh_prim_paths[0]: _OracleMeshHandle(mesh)}
    s._initialize_rays_impl()
    return s


def build_terrain(cfg, spec):
    t = _new(TerrainImporter)
    t.cfg, t.device = cfg, "cpu"
    cur = spec.get("curriculum")
    t.env_origins = torch.as_tensor(spec["env_origins"], dtype=torch.float32).clone()
    if cur:
        t.terrain_origins = torch.as_tensor(cur["terrain_origins"], dtype=torch.float32).clone()
        t.terrain_levels = torch.as_tensor(cur["terrain_levels"], dtype=torch.int64).clone()
        t.terrain_types = torch.as_tensor(cur["terrain_types"], dtype=torch.int64).clone()
        t.max_terrain_level = t.terrain_origins.shape[0]
    else:
        t.terrain_origins = None
    return t


def build_scene(scene_cfg, spec, replay, tmpdir):
    sc = _new(InteractiveScene)
    sc.cfg = scene_cfg
    sc._articulations, sc._deformable_objects, sc._rigid_objects, sc._rigid_object_collections = {}, {}, {}, {}
    sc._sensors, sc._extras, sc._terrain = {}, {}, None
    sc._default_env_origins = torch.as_tensor(spec["env_origins"], dtype=torch.float32).clone()
    robot = build_articulation(scene_cfg.robot, spec, replay, tmpdir)
    sc._articulations["robot"] = robot
    if getattr(scene_cfg, "terrain", None) is not None:
        sc._terrain = build_terrain(scene_cfg.terrain, spec)
    if getattr(scene_cfg, "contact_forces", None) is not None:
        sc._sensors["contact_forces"] = build_contact_sensor(scene_cfg.contact_forces, spec, replay)
    if getattr(scene_cfg, "height_scanner", None) is not None and spec.get("raycaster"):
        sc._sensors["height_scanner"] = build_ray_caster(scene_cfg.height_scanner, spec, robot.root_physx_view)
    return sc


REF_TASK_CFG = {
    "Isaac-Cartpole-v0": ("isaaclab_tasks.manager_based.classic.cartpole.cartpole_env_cfg", "CartpoleEnvCfg"),
    "Isaac-Velocity-Flat-Anymal-C-v0": ("isaaclab_tasks.manager_based.locomotion.velocity.config.anymal_c.flat_env_cfg",
                                        "AnymalCFlatEnvCfg"),
    "Isaac-Velocity-Rough-Anymal-C-v0": ("isaaclab_tasks.manager_based.locomotion.velocity.config.anymal_c.rough_env_cfg",
                                         "AnymalCRoughEnvCfg"),
    "Isaac-Velocity-Rough-Unitree-Go1-v0": ("isaaclab_tasks.manager_based.locomotion.velocity.config.go1.rough_env_cfg",
                                            "UnitreeGo1RoughEnvCfg"),
    "Isaac-Velocity-Rough-H1-v0": ("isaaclab_tasks.manager_based.locomotion.velocity.config.h1.rough_env_cfg",
                                   "H1RoughEnvCfg"),
}


COVERAGE_TASK = "B200-Opcode-Coverage-v0"


def coverage_reference_cfg():
    """The reference-side twin of isaaclab_b200.tasks.coverage.OpcodeCoverageEnvCfg: the reference's flat Anymal-C cfg with
    the reference's OWN IdealPDActuatorCfg / DCMotorCfg classes and mdp functions plugged in, from the same data tables."""
    import importlib
    from isaaclab_b200.tasks import coverage as cov
    from isaaclab.actuators import DCMotorCfg, DelayedPDActuatorCfg, IdealPDActuatorCfg, RemotizedPDActuatorCfg
    from isaaclab.utils import modifiers as rmod
    import isaaclab.envs.mdp as rmdp
    from isaaclab.managers import ObservationTermCfg, RewardTermCfg, SceneEntityCfg, TerminationTermCfg
    from isaaclab.utils.noise import AdditiveUniformNoiseCfg as Unoise
    flat = importlib.import_module(REF_TASK_CFG["Isaac-Velocity-Flat-Anymal-C-v0"][0])
    cfg = flat.AnymalCFlatEnvCfg()
    acts = {}
    for name, a in cov.ACTUATORS.items():
        a = dict(a)
        cls = {"ideal_pd": IdealPDActuatorCfg, "dc_motor": DCMotorCfg, "delayed_pd": DelayedPDActuatorCfg,
               "remotized_pd": RemotizedPDActuatorCfg}[a.pop("kind")]
        acts[name] = cls(**a)
    cfg.scene.robot = cfg.scene.robot.replace(actuators=acts)
    cfg.actions.joint_pos.clip = dict(cov.ACTION_CLIP)
    for name, r in cov.REWARDS.items():
        params = {}
        for k, v in r["params"].items():
            if k == "sensor_cfg":
                v = SceneEntityCfg(v[0], body_names=v[1])
            elif k == "asset_cfg":
                v = SceneEntityCfg(v[0], joint_names=v[1])
            params[k] = v
        setattr(cfg.rewards, "cov_" + name, RewardTermCfg(func=getattr(rmdp, name), weight=r["weight"], params=params))
    for name, t in cov.TERMINATIONS.items():
        setattr(cfg.terminations, "cov_" + name, TerminationTermCfg(func=getattr(rmdp, name), params=dict(t["params"])))
    for name, (fn, nz, clip, scale, hist, mods) in cov.OBSERVATIONS.items():
        mods = [rmod.ModifierCfg(func=getattr(rmod, m), params=dict(pr)) for m, pr in mods] if mods else None
        setattr(cfg.observations.policy, name,
                ObservationTermCfg(func=getattr(rmdp, fn), noise=Unoise(n_min=nz[0], n_max=nz[1]) if nz else None, clip=clip,
                                   scale=scale, history_length=hist, modifiers=mods))
    from isaaclab.managers import EventTermCfg
    for name, m in cov.EVENT_GATES.items():
        getattr(cfg.events, name).min_step_count_between_reset = m
    cfg.events.push_all = EventTermCfg(func=rmdp.push_by_setting_velocity, mode="interval", is_global_time=True,
                                       interval_range_s=cov.GLOBAL_PUSH["interval_range_s"],
                                       params={"velocity_range": dict(cov.GLOBAL_PUSH["velocity_range"])})
    return cfg


def hard_joint_limits(task, spec):
    """Joint position limits the reference would read from USD (stated in isaaclab_b200.assets)."""
    from isaaclab_b200.tasks import TASKS
    from isaaclab_b200.utils.string import resolve_matching_names_values
    robot = TASKS[task]().scene.robot
    idx, _, vals = resolve_matching_names_values(robot.joint_pos_limits, spec["joint_names"])
    lim = torch.zeros(spec["J"], 2)
    lim[idx] = torch.tensor(vals, dtype=torch.float32)
    return lim


def build_reference_env(task, spec, slabs, init, tmpdir, wrap=False):
    import importlib
    if task == COVERAGE_TASK:
        cfg = coverage_reference_cfg()
    else:
        mod, cls = REF_TASK_CFG[task]
        cfg = getattr(importlib.import_module(mod), cls)()
    cfg.scene.num_envs = spec["N"]
    # startup-mode events randomise PhysX materials / masses: PhysX-side, outside the path
    for name in list(vars(cfg.events)):
        term = getattr(cfg.events, name)
        if term is not None and getattr(term, "mode", None) == "startup":
            setattr(cfg.events, name, None)
    # debug visualisation draws USD markers: Kit-only
    for name in (vars(cfg.commands) if hasattr(cfg.commands, "__dict__") else ()):
        term = getattr(cfg.commands, name)
        if hasattr(term, "debug_vis"):
            term.debug_vis = False
    replay = Replay(slabs, wrap=wrap)
    env = _new(ManagerBasedRLEnv)
    env._is_closed = True
    env.cfg = cfg
    env.sim = ReplaySim(replay, cfg.sim.dt)
    env.scene = build_scene(cfg.scene, spec, replay, tmpdir)
    env.extras = {}
    env._sim_step_counter = 0
    env.common_step_counter = 0
    env.episode_length_buf = init["episode_length_buf"].clone()
    # manager construction order of ManagerBasedRLEnv.load_managers / ManagerBasedEnv.load_managers
    env.command_manager = CommandManager(cfg.commands, env)
    env.event_manager = EventManager(cfg.events, env)
    env.recorder_manager = RecorderManager(cfg.recorders, env)
    env.action_manager = ActionManager(cfg.actions, env)
    env.observation_manager = ObservationManager(cfg.observations, env)
    env.termination_manager = TerminationManager(cfg.terminations, env)
    env.reward_manager = RewardManager(cfg.rewards, env)
    env.curriculum_manager = CurriculumManager(cfg.curriculum, env)
    # initial counters shared with the oracle / CUDA env (tests/mdp_common.init_*_state)
    if spec.get("command"):
        term = env.command_manager.get_term(spec["command"]["name"])
        term.time_left[:] = init["cmd_time_left"]
    from isaaclab_b200.sim.synthetic import initial_delay_lags
    for (name, actuator), aspec in zip(env.scene["robot"].actuators.items(), spec["actuators"]):
        if aspec["kind"] in ("delayed_pd", "remotized_pd"):
            lags = initial_delay_lags(aspec, init).to(torch.int)
            for buf in (actuator.positions_delay_buffer, actuator.velocities_delay_buffer, actuator.efforts_delay_buffer):
                buf.set_time_lag(lags)
    for tl in env.event_manager._interval_term_time_left:
        tl[:] = init["interval_time_left"] if tl.numel() > 1 else init.get("interval_global_time_left", 0.03)   # [1]: is_global_time
    return env, replay


def snapshot(env, spec, replay):
    """State of the reference after one step, named as in oracle.mdp_env.OracleEnv."""
    robot = env.scene["robot"]
    d = robot.data
    out = dict(
        episode_length_buf=env.episode_length_buf.clone(), reset_buf=env.reset_buf.clone(),
        term_dones=torch.stack([env.termination_manager._term_dones[n] for n in env.termination_manager._term_names]).clone(),
        episode_sums=torch.stack([env.reward_manager._episode_sums[n] for n in env.reward_manager._term_names]).clone(),
        step_reward=env.reward_manager._step_reward.clone(),
        action=env.action_manager.action.clone(), prev_action=env.action_manager.prev_action.clone(),
        joint_pos_target=d.joint_pos_target.clone(), computed_torque=d.computed_torque.clone(),
        applied_torque=d.applied_torque.clone(), joint_acc=d.joint_acc.clone(),
        root_state_w=replay.root.clone(), joint_pos=replay.joint_pos.clone(), joint_vel=replay.joint_vel.clone(),
        env_origins=env.scene.env_origins.clone(),
    )
    act_state = []
    for a in robot.actuators.values():
        st = {}
        if hasattr(a, "sea_hidden_state"):
            st = {"h": a.sea_hidden_state.clone(), "c": a.sea_cell_state.clone()}
        elif hasattr(a, "_joint_pos_error_history"):
            st = {"pos_hist": a._joint_pos_error_history.clone(), "vel_hist": a._joint_vel_history.clone()}
        act_state.append(st)
    out["act_state"] = act_state
    if spec.get("command"):
        t = env.command_manager.get_term(spec["command"]["name"])
        out.update(cmd_vel_b=t.vel_command_b.clone(), cmd_heading_target=t.heading_target.clone(),
                   cmd_is_heading=t.is_heading_env.clone(), cmd_is_standing=t.is_standing_env.clone(),
                   cmd_time_left=t.time_left.clone(), cmd_counter=t.command_counter.clone(),
                   cmd_metric_xy=t.metrics["error_vel_xy"].clone(), cmd_metric_yaw=t.metrics["error_vel_yaw"].clone())
    if spec.get("contact"):
        c = env.scene.sensors["contact_forces"]._data
        out.update(cs_hist=c.net_forces_w_history.clone(), cs_last_air=c.last_air_time.clone(),
                   cs_cur_air=c.current_air_time.clone(), cs_last_contact=c.last_contact_time.clone(),
                   cs_cur_contact=c.current_contact_time.clone())
    if spec.get("raycaster"):
        r = env.scene.sensors["height_scanner"]
        out.update(rc_hits=r._data.ray_hits_w.clone(), rc_pos_w=r._data.pos_w.clone(), rc_drift=r.drift.clone())
    if spec.get("curriculum"):
        out.update(terrain_levels=env.scene.terrain.terrain_levels.clone())
    out["interval_time_left"] = [t.clone() for t in env.event_manager._interval_term_time_left]
    log = {}
    for k, v in env.extras.get("log", {}).items():
        log[k] = float(v)
    out["log"] = log
    return out


def generate(task, N, steps, slab_seed=1234, init_seed=5, action_seed=3, draw_seed=99):
    spec = mdp_common.make_spec(task, N)
    slabs = generate_slabs(spec, steps + 1, seed=slab_seed)
    init = initial_env_state(spec, seed=init_seed)
    init["episode_length_buf"][::7] = spec["max_episode_length"] - 2
    if task == COVERAGE_TASK:   # some bases below the minimum height (the synthetic z never gets there on its own)
        slabs["root_state_w"][:, ::9, 2] -= 0.25
    slabs_in = {k: v.clone() for k, v in slabs.items()}
    spec["joint_pos_limits"] = hard_joint_limits(task, spec)
    log = DrawLog(draw_seed)
    fill_gen = torch.Generator().manual_seed(draw_seed + 1)
    filler = lambda n, m: torch.rand(n, m, generator=fill_gen)  # noqa: E731
    g = torch.Generator().manual_seed(action_seed)
    ref_ray_caster.raycast_mesh = _raycast_mesh_stand_in
    out_steps = []
    with tempfile.TemporaryDirectory() as tmp:
        env, replay = build_reference_env(task, spec, slabs, init, tmp)
        log.install()
        try:
            for k in range(steps):
                action = torch.randn(N, spec["A"], generator=g)
                log.enabled = True
                obs, rew, terminated, time_outs, extras = env.step(action.clone())
                log.enabled = False
                u_r, u_s, u_o, unused = layout_uniforms(log.take(), spec, N, filler)
                st = snapshot(env, spec, replay)
                st.update(action_in=action, u_reset=u_r, u_step=u_s, u_obs=u_o, obs=obs["policy"].clone(), reward=rew.clone(),
                          terminated=terminated.clone(), time_outs=time_outs.clone(), unused_draws=sorted(set(unused)))
                out_steps.append(st)
        finally:
            log.uninstall()
    return dict(task=task, N=N, steps=steps, slabs=slabs_in, init=init, out=out_steps,
                seeds=dict(slab=slab_seed, init=init_seed, action=action_seed, draw=draw_seed),
                reference="IsaacLab 2.1.0 ManagerBasedRLEnv.step (unmodified) over replayed PhysX views; see gen_mdp.py")


SIZES = {COVERAGE_TASK: (48, 8), "Isaac-Cartpole-v0": (64, 8), "Isaac-Velocity-Flat-Anymal-C-v0": (40, 8), "Isaac-Velocity-Rough-Anymal-C-v0": (36, 6),
         "Isaac-Velocity-Rough-Unitree-Go1-v0": (36, 6), "Isaac-Velocity-Rough-H1-v0": (36, 6)}


def fixture_path(task):
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "mdp_" + task.replace("Isaac-", "").replace("-", "_").lower() + ".pt")


def main(tasks=None):
    for task in tasks or (list(REF_TASK_CFG) + [COVERAGE_TASK]):
        N, steps = SIZES[task]
        fx = generate(task, N, steps)
        n_reset = sum(int(s["reset_buf"].sum()) for s in fx["out"])
        torch.save(fx, fixture_path(task))
        print(f"{task}: N={N} steps={steps} resets={n_reset} unused draws={sorted({u for s in fx['out'] for u in s['unused_draws']})}"
              f" -> {os.path.basename(fixture_path(task))} ({os.path.getsize(fixture_path(task)) / 1e6:.2f} MB)")


if __name__ == "__main__":
    main(sys.argv[1:] or None)
