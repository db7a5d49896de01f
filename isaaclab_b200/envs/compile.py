"""Lower reference-style manager cfgs to the neutral env `spec` the fused kernels (and the test oracle) consume.

What the reference does at manager construction -- resolve `SceneEntityCfg` names to joint/body ids
(LAB/managers/manager_base.py:374-390), check term parameters, resolve actuator groups and per-joint gains
(LAB/assets/articulation/articulation.py:1290-1400), build the ray pattern, generate the terrain -- happens here once,
on the host.  Errors are raised here (construction time), never at step time, as in the reference
(manager_base.py:303-351).
"""
from __future__ import annotations

import math

import numpy as np

from .. import terrains as terrain_mod
from ..assets import ArticulationCfg, synthetic_lstm_net, synthetic_mlp_net, load_lstm_net
from ..managers import (ConstantNoiseCfg, GaussianNoiseCfg, ManagerTermBaseCfg, ObservationGroupCfg, SceneEntityCfg,
                        UniformNoiseCfg)
from ..sensors.ray_caster import grid_pattern
from ..utils.string import resolve_matching_names, resolve_matching_names_values
from . import mdp
from .cfg import ManagerBasedRLEnvCfg


class _Entity:
    """find_joints / find_bodies over a name list (what SceneEntityCfg.resolve needs from a scene entity)."""

    def __init__(self, joint_names=None, body_names=None):
        self.joint_names, self.body_names = joint_names, body_names
        self.num_joints = len(joint_names) if joint_names else 0
        self.num_bodies = len(body_names) if body_names else 0

    def find_joints(self, keys, preserve_order=False):
        return resolve_matching_names(keys, self.joint_names, preserve_order)

    def find_bodies(self, keys, preserve_order=False):
        return resolve_matching_names(keys, self.body_names, preserve_order)


def _terms(container):
    if container is None:
        return []
    if hasattr(container, "terms"):
        return container.terms()
    return [(k, v) for k, v in vars(container).items() if v is not None and not k.startswith("_")]


def _ids(sel, n):
    return list(range(n)) if isinstance(sel, slice) else [int(i) for i in sel]


def _per_joint(val, names, default=None):
    """float | dict(regex -> value) | None -> list over `names` (ActuatorBase._parse_joint_parameter)."""
    if val is None:
        return default
    if isinstance(val, (int, float)):
        return [float(val)] * len(names)
    idx, _, vals = resolve_matching_names_values(val, names)
    out = [0.0] * len(names)
    for i, v in zip(idx, vals):
        out[i] = float(v)
    return out


def _resolve_func(name, term_cfg, min_argc=1):
    """ManagerBase._resolve_common_term_cfg (LAB/managers/manager_base.py:278-351): "module:attr" strings -> callables,
    class terms must derive from ManagerTermBase, and the term's signature must be matched by `params` -- the same
    exceptions with the same messages, raised at construction time."""
    import inspect
    from ..managers import ManagerTermBase
    if not isinstance(term_cfg, ManagerTermBaseCfg):
        raise TypeError(f"Configuration for the term '{name}' is not of type ManagerTermBaseCfg."
                        f" Received: '{type(term_cfg)}'.")
    func = term_cfg.func
    if isinstance(func, str):  # string_to_callable, "module:attr" (manager_base.py:317-318)
        import importlib
        try:
            mod, attr = func.split(":")
            func = getattr(importlib.import_module(mod), attr)
        except Exception as e:
            raise ValueError(f"Could not resolve the input string '{func}' into callable object. Received the error:\n {e}.")
        term_cfg.func = func
    if not callable(func):
        raise AttributeError(f"The term '{name}' is not callable. Received: {func}")
    if inspect.isclass(func):
        if not issubclass(func, ManagerTermBase):
            raise TypeError(f"Configuration for the term '{name}' is not of type ManagerTermBase."
                            f" Received: '{type(func)}'.")
        func_static = func.__call__
        min_argc += 1  # 'self'
    else:
        func_static = func
    if hasattr(func, "_b200_op"):   # fused identities accept any params; each opcode checks its own in FusedEnv._term
        return func
    term_params = list(term_cfg.params.keys())
    args = inspect.signature(func_static).parameters
    with_defaults = [a for a in args if args[a].default is not inspect.Parameter.empty]
    without_defaults = [a for a in args if args[a].default is inspect.Parameter.empty]
    args = without_defaults + with_defaults
    if len(args) > min_argc:
        if set(args[min_argc:]) != set(term_params + with_defaults):
            raise ValueError(f"The term '{name}' expects mandatory parameters: {without_defaults[min_argc:]}"
                             f" and optional parameters: {with_defaults}, but received: {term_params}.")
    return func


# modules whose functions ARE the fused term library under the reference's own names: a cfg that passes the reference's
# function objects (isaaclab.envs.mdp.*, the locomotion-velocity and cartpole task mdp packages) compiles to the same
# opcodes as one that passes isaaclab_b200.envs.mdp identities
_REFERENCE_MDP_MODULES = ("isaaclab.envs.mdp.", "isaaclab_tasks.manager_based.locomotion.velocity.mdp.",
                          "isaaclab_tasks.manager_based.classic.cartpole.mdp.")


def _fused(term_cfg, name, kinds, min_argc=1):
    """The fused identity behind a term's `func`, or None when the term has to be evaluated in Python every step
    (`func(env, **params) -> Tensor`, LAB/managers/manager_term_cfg.py:29-52) and enters the kernels as an EXTERNAL column."""
    func = _resolve_func(name, term_cfg, min_argc)
    if hasattr(func, "_b200_op"):
        if func._b200_kind not in kinds:
            raise ValueError(f"term '{name}': {func._b200_op} is a {func._b200_kind} term, expected one of {kinds}")
        return func
    modname = (getattr(func, "__module__", "") or "") + "."
    if modname.startswith(_REFERENCE_MDP_MODULES):
        ident = mdp._REGISTRY.get(getattr(func, "__name__", ""))
        if ident is not None and ident._b200_kind in kinds:
            return ident
    return None


def _modifier_kind(func):
    """"scale" | "clip" | "bias" for the three function modifiers (ours or the reference's own), else None."""
    kind = getattr(func, "_b200_modifier", None)
    if kind is None:
        mod = (getattr(func, "__module__", "") or "")
        if mod.startswith("isaaclab.utils.modifiers") and getattr(func, "__name__", "") in ("scale", "clip", "bias"):
            kind = func.__name__
    return kind


def _compile_modifiers(name, mods):
    """`ObservationTermCfg.modifiers` -> [("scale", m) | ("bias", v) | ("clip", lo, hi)] (LAB/utils/modifiers/modifier.py:23-66).
    Class modifiers (DigitalFilter, Integrator) keep per-env state in Python objects and are not on the fused path."""
    out = []
    for m in mods:
        kind = _modifier_kind(m.func)
        if kind == "scale":
            out.append(("scale", float(m.params["multiplier"])))
        elif kind == "bias":
            out.append(("bias", float(m.params["value"])))
        elif kind == "clip":
            lo, hi = m.params["bounds"]
            out.append(("clip", -math.inf if lo is None else float(lo), math.inf if hi is None else float(hi)))
        else:
            raise NotImplementedError(f"observation term '{name}': modifier {m.func} (only scale / clip / bias are fused)")
    if len(out) > 3 or sum(len(m) - 1 for m in out) > 4:
        raise NotImplementedError(f"observation term '{name}': at most 3 modifiers with 4 parameters in total")
    return out


def compile_env_cfg(cfg: ManagerBasedRLEnvCfg, num_envs: int | None = None, terrain: "terrain_mod.TerrainGenerator" = None,
                    seed: int = 0) -> dict:
    robot: ArticulationCfg = cfg.scene.robot
    N = int(num_envs if num_envs is not None else cfg.scene.num_envs)
    jn, bn = list(robot.joint_names), list(robot.body_names)
    J, B = len(jn), len(bn)
    scene = {"robot": _Entity(jn, bn)}
    if cfg.scene.contact_forces is not None:
        scene["contact_forces"] = _Entity(None, bn)  # prim_path ".../Robot/.*": one sensor body per robot body
    if cfg.scene.height_scanner is not None:
        scene["height_scanner"] = _Entity()
    physics_dt = float(cfg.sim.dt)
    step_dt = physics_dt * cfg.decimation
    spec = {"N": N, "J": J, "A": 0, "B": B, "R": 0, "decimation": int(cfg.decimation), "physics_dt": physics_dt,
            "step_dt": step_dt, "max_episode_length_s": float(cfg.episode_length_s),
            "max_episode_length": int(math.ceil(cfg.episode_length_s / step_dt)),  # manager_based_rl_env.py:101-103
            "joint_names": jn, "body_names": bn}
    # --- articulation defaults (articulation.py:1279-1300)
    spec["default_root_state"] = [*robot.init_pos, *robot.init_rot, *robot.init_lin_vel, *robot.init_ang_vel]
    spec["default_joint_pos"] = _per_joint(robot.joint_pos, jn)
    spec["default_joint_vel"] = _per_joint(robot.joint_vel, jn)
    lim = np.zeros((J, 2), dtype=np.float32)
    idx, _, vals = resolve_matching_names_values(robot.joint_pos_limits, jn)
    for i, v in zip(idx, vals):
        lim[i] = v
    mean, rng_ = (lim[:, 0] + lim[:, 1]) / 2, lim[:, 1] - lim[:, 0]  # articulation.py:1260-1266
    soft = np.stack([mean - 0.5 * rng_ * robot.soft_joint_pos_limit_factor,
                     mean + 0.5 * rng_ * robot.soft_joint_pos_limit_factor], axis=1).astype(np.float32)
    spec["soft_joint_pos_limits"] = soft
    soft_vel = np.zeros(J, dtype=np.float32)
    # --- actuators (articulation.py:1320-1400)
    acts = []
    for name, ac in robot.actuators.items():
        ids, names = resolve_matching_names(ac.joint_names_expr, jn)
        a = {"name": name, "kind": ac.kind, "joint_ids": ids,
             "stiffness": _per_joint(ac.stiffness, names, [0.0] * len(ids)),
             "damping": _per_joint(ac.damping, names, [0.0] * len(ids)),
             "effort_limit": _per_joint(ac.effort_limit, names, [float("inf")] * len(ids)),
             "velocity_limit": _per_joint(ac.velocity_limit, names, [float("inf")] * len(ids)),
             "saturation_effort": getattr(ac, "saturation_effort", None)}
        if ac.kind in ("delayed_pd", "remotized_pd"):
            a["min_delay"], a["max_delay"] = int(ac.min_delay), int(ac.max_delay)
            if not 0 <= a["min_delay"] <= a["max_delay"]:
                raise ValueError(f"actuator '{name}': need 0 <= min_delay <= max_delay")
        if ac.kind == "remotized_pd":   # actuator_pd.py:372-385: box limits removed, torque limit from the lookup table
            a["effort_limit"] = [float("inf")] * len(ids)
            a["velocity_limit"] = [float("inf")] * len(ids)
            tab = np.asarray(ac.joint_parameter_lookup, dtype=np.float32).reshape(-1, 3)
            if tab.shape[0] == 0:
                raise ValueError("Input tensor x is empty!")                              # linear_interpolation.py:46
            if np.any(tab[1:, 0] < tab[:-1, 0]):
                raise ValueError("Input tensor x is not sorted in ascending order!")     # linear_interpolation.py:50
            a["lookup_angle"], a["lookup_torque"] = tab[:, 0].copy(), tab[:, 2].copy()
        if ac.kind == "lstm":
            nf = ac.network_file
            a["lstm"] = nf if isinstance(nf, dict) else (load_lstm_net(nf) if nf else synthetic_lstm_net(seed=seed))
        elif ac.kind == "mlp":
            nf = ac.network_file
            a["mlp"] = nf if isinstance(nf, dict) else synthetic_mlp_net(ac, seed=seed)
        soft_vel[ids] = a["velocity_limit"]  # articulation.py:1466
        acts.append(a)
    spec["actuators"] = acts
    spec["soft_joint_vel_limits"] = soft_vel
    # --- actions (one joint action term)
    aterms = _terms(cfg.actions)
    if len(aterms) != 1:
        raise NotImplementedError("the fused path supports exactly one joint action term")
    aname, ac = aterms[0]
    ids, names = resolve_matching_names(ac.joint_names, jn, ac.preserve_order)
    A = len(ids)
    scale = ac.scale if isinstance(ac.scale, (int, float)) else None
    if scale is None:
        scale = np.ones(A, dtype=np.float32)
        i2, _, v2 = resolve_matching_names_values(ac.scale, names)
        scale[i2] = v2
    offset = ac.offset if isinstance(ac.offset, (int, float)) else None
    if offset is None:
        offset = np.zeros(A, dtype=np.float32)
        i2, _, v2 = resolve_matching_names_values(ac.offset, names)
        offset[i2] = v2
    kind = "joint_position" if ac.class_type is mdp.JointPositionAction else "joint_effort"
    if kind == "joint_position" and getattr(ac, "use_default_offset", False):
        offset = np.asarray(spec["default_joint_pos"], dtype=np.float32)[ids]
    clip = None
    if ac.clip is not None:
        clip = np.tile(np.array([[-np.inf, np.inf]], dtype=np.float32), (A, 1))
        i2, _, v2 = resolve_matching_names_values(ac.clip, names)
        clip[i2] = v2
    spec["A"] = A
    spec["action"] = {"name": aname, "kind": kind, "joint_ids": ids, "scale": scale, "offset": offset, "clip": clip}
    # --- sensors
    cs = cfg.scene.contact_forces
    if cs is not None:
        spec["contact"] = {"history_length": int(cs.history_length), "track_air_time": bool(cs.track_air_time),
                           "force_threshold": float(cs.force_threshold), "update_period": float(cs.update_period)}
    # --- terrain + env origins
    tcfg = cfg.scene.terrain
    rng = np.random.default_rng(seed)
    if tcfg is not None and tcfg.terrain_type == "generator":
        if terrain is None:
            terrain = terrain_mod.TerrainGenerator(tcfg.terrain_generator)
        origins, levels, types = terrain_mod.compute_env_origins_curriculum(N, terrain.terrain_origins,
                                                                            tcfg.max_init_terrain_level, rng)
        spec["env_origins"] = origins
        terrain_info = {"terrain_origins": terrain.terrain_origins, "terrain_levels": levels, "terrain_types": types,
                        "terrain_size_x": float(tcfg.terrain_generator.size[0])}
    else:
        spec["env_origins"] = terrain_mod.compute_env_origins_grid(N, cfg.scene.env_spacing)
        terrain_info = None
    hs = cfg.scene.height_scanner
    if hs is not None:
        pc = hs.pattern_cfg
        if type(pc).__name__ in ("LidarPatternCfg", "BpearlPatternCfg"):   # sensors/patterns.py: general (non-vertical) rays
            st_, dr_ = pc.func(pc, "cpu")
            starts, dirs = st_.numpy().astype(np.float32), dr_.numpy().astype(np.float32)
        elif type(pc).__name__ == "PinholeCameraPatternCfg":
            raise NotImplementedError("the pinhole pattern belongs to RayCasterCamera (per-env intrinsics), not to RayCaster")
        else:
            starts, dirs = grid_pattern(pc.size, pc.resolution, pc.direction, pc.ordering)
        if tuple(hs.offset.rot) != (1.0, 0.0, 0.0, 0.0):
            raise NotImplementedError("ray caster offset rotation")
        starts = starts + np.asarray(hs.offset.pos, dtype=np.float32)  # ray_caster.py:207-208
        if terrain is not None:
            V, F = terrain.terrain_mesh
        else:  # ground plane of the `plane` terrain type
            V = np.array([[-1e3, -1e3, 0], [1e3, -1e3, 0], [1e3, 1e3, 0], [-1e3, 1e3, 0]], dtype=np.float32)
            F = np.array([[0, 1, 2], [0, 2, 3]], dtype=np.int32)
        spec["R"] = int(starts.shape[0])
        spec["raycaster"] = {"ray_starts": starts, "ray_dirs": dirs, "update_period": float(hs.update_period),
                             "max_distance": float(hs.max_distance), "attach_yaw_only": bool(hs.attach_yaw_only),
                             "drift_range": tuple(hs.drift_range), "mesh_vertices": V, "mesh_faces": F}

    def resolve_params(name, params):
        out = {}
        for k, v in params.items():
            if isinstance(v, SceneEntityCfg):
                v.resolve(scene)
                if v.name == "robot" and k == "asset_cfg":
                    if v.joint_names is not None:
                        out["joint_ids"] = _ids(v.joint_ids, J)
                    if v.body_names is not None:
                        out["asset_body_ids"] = _ids(v.body_ids, B)
                elif v.name == "contact_forces":
                    out["body_ids"] = _ids(v.body_ids, B)
                elif v.name == "height_scanner":
                    pass
                else:
                    raise NotImplementedError(f"term '{name}': scene entity '{v.name}'")
            elif k == "command_name":
                if cfg.commands is None or v not in dict(_terms(cfg.commands)):
                    raise ValueError(f"term '{name}': command '{v}' does not exist")
            else:
                out[k] = v
        return out

    def external(name, t):
        """A term evaluated in Python every step as `func(env, **params)` (manager_term_cfg.py:29-52): SceneEntityCfg
        params are resolved in place and handed to the function as objects, as the reference does
        (manager_base.py:374-390); class terms are instantiated when the env is built (manager_base.py:393-395)."""
        import inspect
        for v in t.params.values():
            if isinstance(v, SceneEntityCfg):
                v.resolve(scene)
        return {"name": name, "op": "external", "func": t.func, "params": dict(t.params), "is_class": inspect.isclass(t.func),
                "cfg": t}

    # --- terminations / rewards
    spec["terminations"] = []
    for name, t in _terms(cfg.terminations):
        f = _fused(t, name, ("term",))
        d = {"op": f._b200_op, **resolve_params(name, t.params)} if f is not None else external(name, t)
        spec["terminations"].append({**d, "name": name, "time_out": bool(t.time_out)})
    spec["rewards"] = []
    for name, t in _terms(cfg.rewards):
        if not isinstance(t.weight, (float, int)):
            raise TypeError(f"Weight for the term '{name}' is not of type float or int. Received: '{type(t.weight)}'.")
        f = _fused(t, name, ("rew",))
        d = {"op": f._b200_op, **resolve_params(name, t.params)} if f is not None else external(name, t)
        spec["rewards"].append({**d, "name": name, "weight": float(t.weight)})
    # --- observations: the policy group (concatenated)
    groups = _terms(cfg.observations)
    if len(groups) != 1:
        raise NotImplementedError("the fused path computes one concatenated observation group")
    gname, grp = groups[0]
    gcfg = getattr(grp, "_group", None) or ObservationGroupCfg()
    if not gcfg.concatenate_terms:
        raise NotImplementedError("concatenate_terms=False")
    oterms = []
    for name, t in _terms(grp):
        if isinstance(t, ObservationGroupCfg):
            continue
        import inspect
        f = _fused(t, name, ("obs",))
        # class modifiers (DigitalFilter, Integrator: modifier.py:69-260) keep per-env state in Python objects: a term that
        # carries one is evaluated in Python (fused identities through their one-term program) and enters as an EXTERNAL column
        class_mods = any(inspect.isclass(m.func) or _modifier_kind(m.func) is None for m in (t.modifiers or ()))   # or any user modifier
        if f is not None and not class_mods:
            d = {"name": name, "op": f._b200_op, **f._b200_defaults, **resolve_params(name, t.params)}
        else:   # width inferred by calling the term once when the env is built (observation_manager.py:420-450)
            d = {**external(name, t), "dim": None}
            if class_mods:
                d["py_modifiers"] = list(t.modifiers)
        # history: the group's setting overrides the terms' (observation_manager.py:411-414)
        hist = gcfg.history_length if gcfg.history_length is not None else t.history_length
        flat = gcfg.flatten_history_dim if gcfg.history_length is not None else t.flatten_history_dim
        if hist:
            if hist < 0:
                raise ValueError(f"observation term '{name}': history_length must be >= 0")
            if not flat:
                raise NotImplementedError(f"observation term '{name}': flatten_history_dim=False in a concatenated group")
            d["history_length"] = int(hist)
        if t.modifiers and not class_mods:
            d["modifiers"] = _compile_modifiers(name, t.modifiers)
        if t.noise is not None and t.noise.operation not in ("add", "scale", "abs"):
            raise ValueError(f"Unknown operation in noise: {t.noise.operation}")   # noise_model.py:41,69,97
        if isinstance(t.noise, UniformNoiseCfg):
            d["noise"] = ("uniform", float(t.noise.n_min), float(t.noise.n_max), t.noise.operation)
        elif isinstance(t.noise, GaussianNoiseCfg):
            d["noise"] = ("gaussian", float(t.noise.mean), float(t.noise.std), t.noise.operation)
        elif isinstance(t.noise, ConstantNoiseCfg):
            d["noise"] = ("constant", float(t.noise.bias), 0.0, t.noise.operation)
        elif t.noise is not None:
            raise NotImplementedError(f"observation term '{name}': {type(t.noise).__name__}")
        d["clip"] = tuple(t.clip) if t.clip is not None else None
        d["scale"] = t.scale
        oterms.append(d)
    spec["observations"] = {"group": gname, "enable_corruption": bool(gcfg.enable_corruption), "terms": oterms}
    # --- commands
    cterms = _terms(cfg.commands)
    if len(cterms) > 1:
        raise NotImplementedError("one command term")
    if cterms:
        cname, c = cterms[0]
        if c.class_type is not mdp.UniformVelocityCommand:
            raise NotImplementedError("command term class")
        if c.heading_command and c.ranges.heading is None:
            raise ValueError("The velocity command has heading commands active (heading_command=True) but the "
                             "`ranges.heading` parameter is set to None.")
        spec["command"] = {"name": cname, "resampling_time_range": tuple(c.resampling_time_range),
                           "rel_standing_envs": c.rel_standing_envs, "rel_heading_envs": c.rel_heading_envs,
                           "heading_command": c.heading_command, "heading_control_stiffness": c.heading_control_stiffness,
                           "ranges": {"lin_vel_x": c.ranges.lin_vel_x, "lin_vel_y": c.ranges.lin_vel_y,
                                      "ang_vel_z": c.ranges.ang_vel_z, "heading": c.ranges.heading}}
    # --- events
    ev = {"reset_joints": [], "interval_push": []}
    for name, t in _terms(cfg.events):
        f = _fused(t, name, ("event",), min_argc=2)
        if f is None:
            raise NotImplementedError(f"event '{name}': user event terms write simulator state from Python; only the "
                                      "event functions of isaaclab_b200.envs.mdp are on the fused path")
        op = f._b200_op
        if t.mode == "startup" or op == "apply_external_force_torque":
            continue  # PhysX property / wrench writers: simulator side, outside the non-physics step
        gate = int(t.min_step_count_between_reset or 0)
        if op == "reset_root_state_uniform" and t.mode == "reset":
            if "reset_root_state" in ev:
                raise NotImplementedError("more than one reset_root_state_uniform term")
            ev["reset_root_state"] = {"pose_range": dict(t.params["pose_range"]),
                                      "velocity_range": dict(t.params["velocity_range"]), "min_step_count": gate}
        elif op in ("reset_joints_by_scale", "reset_joints_by_offset") and t.mode == "reset":
            ev["reset_joints"].append({"mode": "scale" if op.endswith("scale") else "offset",
                                       "position_range": tuple(t.params["position_range"]),
                                       "velocity_range": tuple(t.params["velocity_range"]), "min_step_count": gate})
        elif op == "push_by_setting_velocity" and t.mode == "interval":
            ev["interval_push"].append({"interval_range_s": tuple(t.interval_range_s), "is_global_time": bool(t.is_global_time),
                                        "velocity_range": dict(t.params["velocity_range"])})
        else:
            raise NotImplementedError(f"event '{name}' ({op}, mode={t.mode})")
    spec["events"] = ev
    # --- curriculum
    for name, t in _terms(cfg.curriculum):
        if _fused(t, name, ("curriculum",), min_argc=2) is None:
            raise NotImplementedError(f"curriculum '{name}': only terrain_levels_vel is on the fused path")
        if terrain_info is None:
            raise ValueError("terrain_levels_vel needs the 'generator' terrain type")
        spec["curriculum"] = {"name": name, **terrain_info}
    return spec
