"""The drop-in boundary on the GPU (Contract 1):

* user (Python) terms -- one lambda per manager: reward, termination, observation -- evaluated every step as
  `func(env, **params)` and fed to the kernels through the EXTERNAL columns, against an env whose cfg uses the fused twins of
  the same terms (`joint_vel_l2`, `root_height_below_minimum`, `joint_pos_rel`): same masks bit for bit, same rewards /
  observations to fp32 summation order;
* a stateful class term (`ManagerTermBase`): constructed once, called every step, `reset(env_ids)` for the reset envs;
* the manager objects' surface (`reset -> dict`, `get_term_cfg` / `set_term_cfg`, `compute_group`, `get_term`, ...);
* a live simulator backend (`LiveSimBackend.physics_step`) against the replay path: identical bits."""
import copy

import pytest
import torch

from isaaclab_b200.envs import mdp
from isaaclab_b200.managers import (ManagerTermBase, ObservationTermCfg, RewardTermCfg, SceneEntityCfg, TerminationTermCfg,
                                    UniformNoiseCfg)
from tests.mdp_common import small_terrain

pytestmark = pytest.mark.gpu

TASK = "Isaac-Velocity-Rough-Anymal-C-v0"
N = 256


def _env(cfg, seed=3, **kw):
    from isaaclab_b200.envs.manager_based_rl_env import DeviceReplaySim, ManagerBasedRLEnv
    from isaaclab_b200.envs.compile import compile_env_cfg
    from isaaclab_b200.sim.synthetic import SyntheticSim, generate_slabs
    cfg.scene.terrain.max_init_terrain_level = 2
    spec = compile_env_cfg(cfg, num_envs=N, terrain=small_terrain(), seed=0)
    slabs = generate_slabs(spec, 8, seed=77)
    slabs["root_state_w"][:, ::9, 2] -= 0.45        # some bases low enough for the height termination
    sim = DeviceReplaySim(SyntheticSim(spec, K=8, device="cuda", slabs=slabs))
    env = ManagerBasedRLEnv(cfg, sim=sim, device="cuda", spec=spec, seed=seed, **kw)
    env.episode_length_buf[::5] = spec["max_episode_length"] - 3
    return env


def _user_joint_vel_l2(env, asset_cfg: SceneEntityCfg = SceneEntityCfg("robot")):
    return torch.sum(torch.square(env.scene[asset_cfg.name].data.joint_vel[:, asset_cfg.joint_ids]), dim=1)


def _user_root_height(env, minimum_height: float):
    return env.scene["robot"].data.root_pos_w[:, 2] < minimum_height


def _user_joint_pos_rel(env):
    d = env.scene["robot"].data
    return d.joint_pos - d.default_joint_pos


def _cfgs():
    from isaaclab_b200.tasks import TASKS
    fused, user = TASKS[TASK](), TASKS[TASK]()
    sel = SceneEntityCfg("robot", joint_names=[".*KFE"])
    fused.rewards.jv = RewardTermCfg(func=mdp.joint_vel_l2, weight=-1e-3, params={"asset_cfg": sel})
    user.rewards.jv = RewardTermCfg(func=_user_joint_vel_l2, weight=-1e-3, params={"asset_cfg": copy.deepcopy(sel)})
    fused.terminations.low = TerminationTermCfg(func=mdp.root_height_below_minimum, params={"minimum_height": 0.3})
    user.terminations.low = TerminationTermCfg(func=_user_root_height, params={"minimum_height": 0.3})
    noise = UniformNoiseCfg(n_min=-0.02, n_max=0.02)
    fused.observations.policy.jp2 = ObservationTermCfg(func=mdp.joint_pos_rel, noise=noise, clip=(-0.6, 0.6), scale=2.0)
    user.observations.policy.jp2 = ObservationTermCfg(func=_user_joint_pos_rel, noise=copy.deepcopy(noise), clip=(-0.6, 0.6), scale=2.0)
    return fused, user


def test_user_terms_match_their_fused_twins():
    cfg_f, cfg_u = _cfgs()
    env_f, env_u = _env(cfg_f), _env(cfg_u)
    assert env_u.fused.n_ext == 2 + 12 and env_f.fused.n_ext == 0
    assert env_u.fused.obs_dim == env_f.fused.obs_dim == 235 + 12
    g = torch.Generator().manual_seed(1)
    (o_f, _), (o_u, _) = env_f.reset(), env_u.reset()
    torch.testing.assert_close(o_u["policy"], o_f["policy"], rtol=1e-6, atol=1e-6)
    n_low = n_reset = 0
    for k in range(6):
        a = torch.randn(N, 12, generator=g).cuda()
        obs_f, rew_f, term_f, to_f, _ = env_f.step(a)
        obs_u, rew_u, term_u, to_u, _ = env_u.step(a)
        torch.cuda.synchronize()
        assert torch.equal(term_f, term_u) and torch.equal(to_f, to_u), f"step {k}: masks differ"
        assert torch.equal(env_f.termination_manager.get_term("low"), env_u.termination_manager.get_term("low"))
        torch.testing.assert_close(rew_u, rew_f, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(env_u.reward_manager._episode_sums["jv"], env_f.reward_manager._episode_sums["jv"], rtol=1e-5, atol=1e-7)
        torch.testing.assert_close(obs_u["policy"], obs_f["policy"], rtol=1e-6, atol=1e-6)
        n_low += int(env_u.termination_manager.get_term("low").sum())
        n_reset += int(env_u.reset_buf.sum())
    assert n_low > 0 and n_reset > n_low, "the horizon did not exercise the user termination / the time-outs"
    # the untouched part of the managers agrees too (resets, commands, curriculum went through the same kernels)
    assert torch.equal(env_f.fused.cmd_counter, env_u.fused.cmd_counter)
    assert torch.equal(env_f.fused.terrain_levels, env_u.fused.terrain_levels)


class _StepsSinceReset(ManagerTermBase):
    """Stateful user term: number of steps since this env's last reset, scaled."""

    def __init__(self, cfg, env):
        super().__init__(cfg, env)
        self.count = torch.zeros(env.num_envs, device=env.device)
        self.resets_seen = 0

    def reset(self, env_ids=None):
        self.count[env_ids] = 0.0
        self.resets_seen += len(env_ids)

    def __call__(self, env, gain: float):
        self.count += 1.0
        return gain * self.count


def test_class_term_is_constructed_once_called_every_step_and_reset_for_reset_envs():
    from isaaclab_b200.tasks import TASKS
    cfg = TASKS[TASK]()
    cfg.rewards.alive_steps = RewardTermCfg(func=_StepsSinceReset, weight=1.0, params={"gain": 0.5})
    env = _env(cfg)
    term = env._class_terms[0]
    assert isinstance(term, _StepsSinceReset) and len(env._class_terms) == 1
    i = env.reward_manager.active_terms.index("alive_steps")
    expect = torch.zeros(N, device="cuda")
    total = 0
    for k in range(5):
        env.step(torch.zeros(N, 12, device="cuda"))
        expect += 1.0
        # _step_reward = weighted value / dt (reward_manager.py:153): the term's value of THIS step, before the reset
        torch.testing.assert_close(env.reward_manager._step_reward[:, i], 0.5 * expect, rtol=1e-6, atol=1e-6)
        expect[env.reset_buf] = 0.0
        total += int(env.reset_buf.sum())
    assert total > 0 and term.resets_seen == total
    torch.testing.assert_close(term.count, expect)


def test_manager_surface():
    from isaaclab_b200.tasks import TASKS
    env = _env(TASKS[TASK](), log_every_step=True)
    env.reset()
    rm, tm, cm, am, om = env.reward_manager, env.termination_manager, env.command_manager, env.action_manager, env.observation_manager
    a = torch.randn(N, 12, device="cuda")
    obs, rew, term, to, extras = env.step(a)
    # ownership: the same tensor objects every step (reward_manager.py:62,157; termination_manager.py:68-69)
    obs2, rew2, term2, to2, _ = env.step(a)
    assert rew2 is rew and term2 is term and to2 is to and obs2["policy"] is obs["policy"]
    assert rm.compute(env.step_dt) is rew and tm.dones is env.reset_buf and tm.compute() is env.reset_buf
    keep = ~env.reset_buf   # ActionManager.reset zeroes the action of the envs reset in this step (action_manager.py:296-316)
    assert torch.equal(am.action[keep], a[keep]) and not am.action[env.reset_buf].any()
    assert am.total_action_dim == 12 and am.get_term("joint_pos").raw_actions is env.fused.raw_actions
    assert om.compute_group("policy") is obs["policy"] and om.compute()["policy"] is obs["policy"]
    assert om.group_obs_dim == {"policy": (235,)} and len(om.active_terms["policy"]) == 8
    assert sum(d[0] for d in om.group_obs_term_dim["policy"]) == 235
    with pytest.raises(ValueError, match="Unable to find the group"):
        om.compute_group("critic")
    assert cm.get_command("base_velocity").shape == (N, 3) and cm.get_term("base_velocity").command is cm.get_command("base_velocity")
    with pytest.raises(ValueError):
        cm.get_command("nope")
    assert set(env.event_manager.available_modes) == {"reset", "interval"} and env.curriculum_manager.active_terms == ["terrain_levels"]
    # extras["log"] is filled per step with the reference's keys; the managers' reset() hands out their share of it
    assert "log" in extras and any(k.startswith("Episode_Reward/") for k in extras["log"])
    log = {}
    for m in (env.curriculum_manager, rm, cm, tm):
        log.update(m.reset(None))
    assert log == env.extras["log"] and "Curriculum/terrain_levels" in log and "Episode_Termination/time_out" in log
    # get_term_cfg / set_term_cfg (reward_manager.py:163-190): a weight change reaches the kernels
    cfg = rm.get_term_cfg("track_lin_vel_xy_exp")
    assert cfg.weight == 1.0
    with pytest.raises(ValueError, match="not found"):
        rm.get_term_cfg("nope")
    i = rm.active_terms.index("track_lin_vel_xy_exp")
    assert float(rm._step_reward[:, i].max()) <= 1.0 + 1e-6      # exp(-err) * weight 1.0
    cfg.weight = 3.0
    rm.set_term_cfg("track_lin_vel_xy_exp", cfg)
    assert rm.get_term_cfg("track_lin_vel_xy_exp").weight == 3.0
    env.step(a)
    torch.cuda.synchronize()
    assert 1.0 < float(rm._step_reward[:, i].max()) <= 3.0 + 1e-6, "the new weight did not reach the step kernel"


class _FakeLiveSim:
    """A 'running simulator' over recorded slabs: `physics_step` moves to the next recorded physics state."""

    def __init__(self, syn):
        self.syn, self.S, self.k, self.s = syn, syn.S, 0, syn.S - 1
        self.steps = 0

    def current_view(self):
        return self.syn.views(self.k)[self.s]

    def last_view(self):
        return self.current_view()

    def physics_step(self, env):
        # a real backend would consume these here (what Articulation.write_data_to_sim hands to PhysX)
        assert env.fused.joint_effort_target_sim.shape == (env.num_envs, env.fused.J)
        self.steps += 1
        self.s += 1
        if self.s == self.S:
            self.k, self.s = self.k + 1, 0


@pytest.mark.parametrize("task", [TASK, "Isaac-Cartpole-v0"])
def test_live_simulator_backend_matches_replay(task):
    from isaaclab_b200.envs.compile import compile_env_cfg
    from isaaclab_b200.envs.manager_based_rl_env import DeviceReplaySim, ManagerBasedRLEnv
    from isaaclab_b200.sim.synthetic import SyntheticSim, generate_slabs
    from isaaclab_b200.tasks import TASKS
    from tests.mdp_common import make_spec
    spec = make_spec(task, N)
    slabs = generate_slabs(spec, 8, seed=5)
    outs = []
    for live in (False, True):
        syn = SyntheticSim(spec, K=8, device="cuda", slabs={k: v.clone() for k, v in slabs.items()})
        sim = _FakeLiveSim(syn) if live else DeviceReplaySim(syn)
        env = ManagerBasedRLEnv(TASKS[task](), sim=sim, device="cuda", spec=copy.deepcopy(spec), seed=3)
        env.episode_length_buf[::5] = spec["max_episode_length"] - 3
        g = torch.Generator().manual_seed(9)
        rec = []
        for _ in range(6):
            obs, rew, term, to, _ = env.step(torch.randn(N, spec["A"], generator=g).cuda())
            rec.append((obs["policy"].clone(), rew.clone(), term.clone(), to.clone()))
        torch.cuda.synchronize()
        if live:
            assert sim.steps == 6 * spec["decimation"]
        outs.append(rec)
    assert any(bool(r[3].any()) for r in outs[0])
    for k, (a, b) in enumerate(zip(*outs)):
        for x, y, name in zip(a, b, ("obs", "reward", "terminated", "time_outs")):
            assert torch.equal(x, y), f"{task} step {k}: {name} differs between the live backend and the replay"


def test_observation_history_through_the_public_env():
    """LAB test_observation_manager.py:357-508 in spirit, through `ManagerBasedRLEnv`: with a group-level history_length the
    observation returned by reset() holds the first row repeated (circular_buffer.py:118-123); every step shifts the window
    by one row for envs that kept running and refills it for the envs the step reset (observation_manager.py:226-230) --
    including the height-scan columns."""
    from isaaclab_b200.tasks import TASKS
    cfg = TASKS[TASK]()
    cfg.observations.policy._group.history_length = 3
    cfg.observations.policy._group.enable_corruption = False
    env = _env(cfg)
    H, dims = 3, [d[0] // 3 for d in env.observation_manager.group_obs_term_dim["policy"]]
    assert sum(dims) == 235 and env.observation_manager.group_obs_dim["policy"] == (705,)

    def rows(obs):   # [N, 705] -> per term [N, H, d]
        out, c = [], 0
        for d in dims:
            out.append(obs[:, c:c + H * d].reshape(N, H, d).clone())
            c += H * d
        return out

    obs, _ = env.reset()
    prev = rows(obs["policy"])
    for r in prev:
        assert torch.equal(r[:, 0], r[:, 2]) and torch.equal(r[:, 1], r[:, 2])
    g = torch.Generator(device="cuda").manual_seed(0)
    n_reset = 0
    for _ in range(5):
        obs, _, term, trunc, _ = env.step(torch.randn(N, 12, device="cuda", generator=g))
        cur = rows(obs["policy"])
        done = term | trunc
        n_reset += int(done.sum())
        for p_, c_ in zip(prev, cur):
            assert torch.equal(c_[~done][:, :2], p_[~done][:, 1:])          # window moved by one row
            assert torch.equal(c_[done][:, 0], c_[done][:, 2]) and torch.equal(c_[done][:, 1], c_[done][:, 2])   # refilled
        assert not torch.equal(cur[3][~done][:, 2], cur[3][~done][:, 1])   # and the newest row is new (joint_pos moves)
        prev = cur
    assert n_reset > 0


def test_recording_hook_reproduces_the_run_it_recorded():
    """sim/recorder.py: a live backend wrapped in `RecordingSim` yields slabs in the replay layout; replaying them through the
    fused (one-launch decimation) path gives bit for bit the observations, rewards and masks of the recorded live run."""
    from isaaclab_b200.envs.manager_based_rl_env import DeviceReplaySim, ManagerBasedRLEnv
    from isaaclab_b200.sim.recorder import RecordingSim
    from isaaclab_b200.sim.synthetic import SyntheticSim, generate_slabs
    from isaaclab_b200.tasks import TASKS
    from tests.mdp_common import make_spec
    spec = make_spec(TASK, N)
    K = 5
    slabs = generate_slabs(spec, K + 1, seed=5)
    rec = RecordingSim(_FakeLiveSim(SyntheticSim(spec, K=K + 1, device="cuda", slabs={k: v.clone() for k, v in slabs.items()})),
                       decimation=spec["decimation"])
    runs = []
    for sim in (rec, None):
        if sim is None:   # second pass: replay what the first pass recorded
            got = rec.slabs()
            assert rec.steps_recorded == K and got["joint_pos"].shape == (K + 1, spec["decimation"], N, spec["J"])
            assert got["root_state_w"].shape == (K + 1, N, 13) and got["net_forces_w"].shape[-2:] == (spec["B"], 3)
            sim = DeviceReplaySim(SyntheticSim(spec, K=K + 1, device="cuda", slabs=got))   # K + 1 slabs: the initial state and K steps
        env = ManagerBasedRLEnv(TASKS[TASK](), sim=sim, device="cuda", spec=copy.deepcopy(spec), seed=3)
        env.episode_length_buf[::5] = spec["max_episode_length"] - 3
        g = torch.Generator().manual_seed(9)
        out = []
        for _ in range(K):
            obs, rew, term, to, _ = env.step(torch.randn(N, spec["A"], generator=g).cuda())
            out.append((obs["policy"].clone(), rew.clone(), term.clone(), to.clone()))
        runs.append(out)
    for k, (a, b) in enumerate(zip(*runs)):
        for x, y, name in zip(a, b, ("obs", "reward", "terminated", "time_outs")):
            assert torch.equal(x, y), f"step {k}: {name} differs between the recorded run and its replay"


def test_class_and_user_modifiers_on_a_fused_term():
    """A term with a class modifier (Integrator) or a user modifier function is evaluated in Python -- the fused identity
    through its one-term program -- modified, and enters the kernels as an EXTERNAL column; noise / clip / scale still run in
    the kernel.  Checked against the same env without the modifier: integral of joint_vel_rel by the trapezoid rule, reset
    to zero for the envs a step resets (observation_manager.py:231-233)."""
    from isaaclab_b200.tasks import TASKS
    from isaaclab_b200.utils import modifiers as MOD
    outs = {}
    for variant in ("plain", "integrated"):
        cfg = TASKS[TASK]()
        cfg.observations.policy._group.enable_corruption = False
        if variant == "integrated":
            cfg.observations.policy.joint_vel.modifiers = [MOD.ModifierCfg(func=lambda data, gain: data * gain, params={"gain": 2.0}),
                                                           MOD.IntegratorCfg(dt=0.5)]
        env = _env(cfg)
        assert ({t["name"]: t["op"] for t in env.spec["observations"]["terms"]}["joint_vel"] ==
                ("external" if variant == "integrated" else "joint_vel_rel"))
        env.reset()
        g = torch.Generator(device="cuda").manual_seed(0)
        rec = []
        for _ in range(5):
            obs, _, term, trunc, _ = env.step(torch.randn(N, 12, device="cuda", generator=g))
            rec.append((obs["policy"][:, 24:36].clone(), (term | trunc).clone()))   # joint_vel columns of the rough policy group
        outs[variant] = rec
    n_done = 0
    for k in range(1, 5):   # step 0's integral also holds the reset-time sample, which the plain run did not record
        (jv, done), (jv_prev, _) = outs["plain"][k], outs["plain"][k - 1]
        (got, done2), (got_prev, _) = outs["integrated"][k], outs["integrated"][k - 1]
        assert torch.equal(done, done2)
        d = done.unsqueeze(1)
        base = torch.where(d, torch.zeros_like(got_prev), got_prev)        # Integrator.reset for the envs this step reset
        x_prev = torch.where(d, torch.zeros_like(jv_prev), 2.0 * jv_prev)
        want = base + (2.0 * jv + x_prev) / 2 * 0.5
        torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-5, msg=lambda m: f"step {k}: {m}")
        n_done += int(done.sum())
    assert n_done > 0
