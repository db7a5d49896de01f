"""tests/golden/task_cfgs.json: the numbers of the reference's own task cfgs (TASKS/.../locomotion/velocity/config/*,
classic/cartpole) that reach the non-physics step -- reward weights and params, termination / observation / event term
names and scalars, action scale, actuator models, command ranges -- read from /root/reference through the stub loader.
Run in the authoring container only:  python tests/golden/gen_task_cfgs.py"""
import importlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness  # noqa: E402

ref_harness.install()

V = "isaaclab_tasks.manager_based.locomotion.velocity.config."
REF = {
    "Isaac-Cartpole-v0": ("isaaclab_tasks.manager_based.classic.cartpole.cartpole_env_cfg", "CartpoleEnvCfg"),
    "Isaac-Velocity-Flat-Anymal-C-v0": (V + "anymal_c.flat_env_cfg", "AnymalCFlatEnvCfg"),
    "Isaac-Velocity-Rough-Anymal-C-v0": (V + "anymal_c.rough_env_cfg", "AnymalCRoughEnvCfg"),
    "Isaac-Velocity-Rough-Anymal-C-Play-v0": (V + "anymal_c.rough_env_cfg", "AnymalCRoughEnvCfg_PLAY"),
    "Isaac-Velocity-Flat-Anymal-B-v0": (V + "anymal_b.flat_env_cfg", "AnymalBFlatEnvCfg"),
    "Isaac-Velocity-Rough-Anymal-B-v0": (V + "anymal_b.rough_env_cfg", "AnymalBRoughEnvCfg"),
    "Isaac-Velocity-Flat-Anymal-D-v0": (V + "anymal_d.flat_env_cfg", "AnymalDFlatEnvCfg"),
    "Isaac-Velocity-Rough-Anymal-D-v0": (V + "anymal_d.rough_env_cfg", "AnymalDRoughEnvCfg"),
    "Isaac-Velocity-Flat-Unitree-A1-v0": (V + "a1.flat_env_cfg", "UnitreeA1FlatEnvCfg"),
    "Isaac-Velocity-Rough-Unitree-A1-v0": (V + "a1.rough_env_cfg", "UnitreeA1RoughEnvCfg"),
    "Isaac-Velocity-Flat-Unitree-Go1-v0": (V + "go1.flat_env_cfg", "UnitreeGo1FlatEnvCfg"),
    "Isaac-Velocity-Rough-Unitree-Go1-v0": (V + "go1.rough_env_cfg", "UnitreeGo1RoughEnvCfg"),
    "Isaac-Velocity-Flat-Unitree-Go2-v0": (V + "go2.flat_env_cfg", "UnitreeGo2FlatEnvCfg"),
    "Isaac-Velocity-Rough-Unitree-Go2-v0": (V + "go2.rough_env_cfg", "UnitreeGo2RoughEnvCfg"),
    "Isaac-Velocity-Rough-Unitree-Go2-Play-v0": (V + "go2.rough_env_cfg", "UnitreeGo2RoughEnvCfg_PLAY"),
    "Isaac-Velocity-Rough-H1-v0": (V + "h1.rough_env_cfg", "H1RoughEnvCfg"),
    "Isaac-Velocity-Flat-H1-v0": (V + "h1.flat_env_cfg", "H1FlatEnvCfg"),
    "Isaac-Velocity-Rough-G1-v0": (V + "g1.rough_env_cfg", "G1RoughEnvCfg"),
    "Isaac-Velocity-Flat-G1-v0": (V + "g1.flat_env_cfg", "G1FlatEnvCfg"),
    "Isaac-Velocity-Rough-Cassie-v0": (V + "cassie.rough_env_cfg", "CassieRoughEnvCfg"),
    "Isaac-Velocity-Flat-Cassie-v0": (V + "cassie.flat_env_cfg", "CassieFlatEnvCfg"),
}


def _plain(v):
    if isinstance(v, (int, float, str, bool)) or v is None:
        return v
    if isinstance(v, (list, tuple)):
        return [_plain(x) for x in v]
    if isinstance(v, dict):
        return {str(k): _plain(x) for k, x in v.items()}
    for attr in ("body_names", "joint_names"):   # SceneEntityCfg
        if hasattr(v, "name") and hasattr(v, attr):
            return {"entity": v.name, "body_names": _plain(v.body_names), "joint_names": _plain(v.joint_names)}
    return type(v).__name__


def _terms(container):
    out = {}
    if container is None:
        return out
    for k, v in vars(container).items():
        if v is None or k.startswith("_") or not hasattr(v, "func"):
            continue
        d = {"func": getattr(v.func, "__name__", str(v.func)), "params": _plain(dict(v.params))}
        for f in ("weight", "time_out", "mode", "interval_range_s", "scale", "clip", "history_length"):
            if hasattr(v, f):
                d[f] = _plain(getattr(v, f))
        if getattr(v, "noise", None) is not None:
            d["noise"] = [type(v.noise).__name__, _plain(getattr(v.noise, "n_min", None)), _plain(getattr(v.noise, "n_max", None))]
        out[k] = d
    return out


def digest(cfg):
    """Works on the reference's cfg objects and on isaaclab_b200's (same attribute names)."""
    d = {"decimation": cfg.decimation, "sim_dt": cfg.sim.dt, "episode_length_s": cfg.episode_length_s,
         "num_envs": cfg.scene.num_envs, "env_spacing": cfg.scene.env_spacing,
         "rewards": _terms(cfg.rewards), "terminations": _terms(cfg.terminations), "events": _terms(cfg.events),
         "curriculum": _terms(cfg.curriculum), "observations": _terms(cfg.observations.policy)}
    pol = cfg.observations.policy
    d["enable_corruption"] = bool(getattr(getattr(pol, "_group", pol), "enable_corruption", False))
    d["actions"] = {k: {"scale": _plain(v.scale), "joint_names": _plain(v.joint_names),
                        "use_default_offset": getattr(v, "use_default_offset", None)}
                    for k, v in vars(cfg.actions).items() if v is not None and hasattr(v, "joint_names")}
    cmd = {}
    for k, v in (vars(cfg.commands).items() if hasattr(cfg.commands, "__dict__") else ()):
        if v is not None and hasattr(v, "ranges"):
            cmd[k] = {"resampling_time_range": _plain(v.resampling_time_range), "rel_standing_envs": v.rel_standing_envs,
                      "rel_heading_envs": v.rel_heading_envs, "heading_command": v.heading_command,
                      "ranges": {r: _plain(getattr(v.ranges, r)) for r in ("lin_vel_x", "lin_vel_y", "ang_vel_z", "heading")}}
    d["commands"] = cmd
    t = getattr(cfg.scene, "terrain", None)
    d["terrain"] = None if t is None else {"type": t.terrain_type, "max_init_terrain_level": t.max_init_terrain_level}
    gen = getattr(t, "terrain_generator", None) if t is not None else None
    if gen is not None:
        d["terrain"].update(num_rows=gen.num_rows, num_cols=gen.num_cols, curriculum=gen.curriculum,
                            sub_terrains={k: {f: _plain(getattr(s, f)) for f in ("proportion", "grid_height_range", "noise_range",
                                                                                "noise_step", "step_height_range", "slope_range")
                                              if hasattr(s, f)} for k, s in gen.sub_terrains.items()})
    hs = getattr(cfg.scene, "height_scanner", None)
    d["height_scanner"] = None if hs is None else {"size": _plain(hs.pattern_cfg.size), "resolution": hs.pattern_cfg.resolution,
                                                     "attach_yaw_only": hs.attach_yaw_only, "offset": _plain(hs.offset.pos)}
    robot = cfg.scene.robot
    init = getattr(robot, "init_state", None)
    d["robot"] = {"init_pos": _plain(init.pos if init is not None else robot.init_pos),
                  "joint_pos": _plain(init.joint_pos if init is not None else robot.joint_pos),
                  "soft_joint_pos_limit_factor": robot.soft_joint_pos_limit_factor,
                  "actuators": {k: {"class": type(a).__name__,
                                    **{f: _plain(getattr(a, f)) for f in ("joint_names_expr", "effort_limit", "velocity_limit",
                                                                         "stiffness", "damping", "saturation_effort", "pos_scale",
                                                                         "vel_scale", "torque_scale", "input_order", "input_idx",
                                                                         "min_delay", "max_delay") if hasattr(a, f)}}
                                for k, a in robot.actuators.items()}}
    return d


if __name__ == "__main__":
    out = {}
    for task, (mod, cls) in REF.items():
        cfg = getattr(importlib.import_module(mod), cls)()
        out[task] = digest(cfg)
        print(task, len(out[task]["rewards"]), "rewards")
    with open(os.path.join(HERE, "task_cfgs.json"), "w") as f:
        json.dump(out, f, indent=1)   # key order = term order (it is the reward summation order)
