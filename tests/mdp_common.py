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


def init_fused_state(env, init):
    env.episode_length_buf.copy_(init["episode_length_buf"])
    if env.spec.get("command"):
        env.cmd_time_left.copy_(init["cmd_time_left"])
    for t in env.interval_time_left:
        t.copy_(init["interval_time_left"])
