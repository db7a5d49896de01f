"""The rsl_rl runner mirror on the GPU: episode bookkeeping against the reference's deque logic
(RSL/runners/on_policy_runner.py:113-132) restated in Python, the VecEnv contract (RSL/env/vec_env.py:13-98), and
`OnPolicyRunner.learn/save/load/get_inference_policy` over the fused env."""
import collections
import os

import pytest
import torch

from tests.mdp_common import make_spec

pytestmark = pytest.mark.gpu


def _deque_reference(rewards_seq, dones_seq, N):
    rew, ln = collections.deque(maxlen=100), collections.deque(maxlen=100)
    cur_r, cur_l = torch.zeros(N), torch.zeros(N)
    for r, d in zip(rewards_seq, dones_seq):
        cur_r += r
        cur_l += 1
        ids = (d > 0).nonzero(as_tuple=False)
        rew.extend(cur_r[ids][:, 0].numpy().tolist())
        ln.extend(cur_l[ids][:, 0].numpy().tolist())
        cur_r[ids] = 0
        cur_l[ids] = 0
    return list(rew), list(ln), cur_r, cur_l


@pytest.mark.parametrize("N,p_done,dtype", [(4096, 0.01, torch.long), (4096, 0.2, torch.uint8), (37, 0.3, torch.bool),
                                            (5000, 0.0, torch.long), (1, 0.5, torch.long)])
def test_episode_stats_match_deque_logic(N, p_done, dtype):
    from isaaclab_b200.rsl_rl import EpisodeStats
    g = torch.Generator().manual_seed(N)
    steps = 30
    rewards = [torch.randn(N, generator=g) for _ in range(steps)]
    dones = [(torch.rand(N, generator=g) < p_done) for _ in range(steps)]
    st = EpisodeStats(N, "cuda")
    for r, d in zip(rewards, dones):
        st.update(r.cuda(), d.to(dtype).cuda())
    rew, ln = st.read()
    e_rew, e_ln, e_cr, e_cl = _deque_reference(rewards, [d.long() for d in dones], N)
    assert rew == e_rew  # same fp32 additions in the same order: bit-exact
    assert ln == e_ln
    assert torch.equal(st.cur_reward_sum.cpu(), e_cr) and torch.equal(st.cur_episode_length.cpu(), e_cl)


def _make_env(task, N):
    from isaaclab_b200.envs.manager_based_rl_env import ManagerBasedRLEnv
    from isaaclab_b200.tasks import TASKS
    spec = make_spec(task, N)
    return ManagerBasedRLEnv(TASKS[task](), spec=spec, device="cuda")


def test_vecenv_wrapper_contract():
    from isaaclab_b200.rsl_rl import RslRlVecEnvWrapper
    env = RslRlVecEnvWrapper(_make_env("Isaac-Velocity-Flat-Anymal-C-v0", 64), clip_actions=1.5)
    assert env.num_envs == 64 and env.num_actions == 12 and env.num_obs == 48 and env.num_privileged_obs == 0
    obs = env.get_observations()
    assert list(obs.keys()) == ["policy"] and obs["policy"].shape == (64, 48)
    env.episode_length_buf = torch.full_like(env.episode_length_buf, int(env.max_episode_length) - 2)
    big = torch.full((64, 12), 9.0, device="cuda")
    saw_timeout = False
    for _ in range(3):
        obs, rew, dones, extras = env.step(big)
        assert obs["policy"].shape == (64, 48) and rew.shape == (64,) and rew.dtype == torch.float32
        assert dones.dtype == torch.long and dones.shape == (64,)
        assert "time_outs" in extras and extras["observations"]["policy"] is obs["policy"]
        base = env.unwrapped
        assert torch.equal(dones.bool(), base.reset_terminated | base.reset_time_outs)
        saw_timeout |= bool(extras["time_outs"].any())
        # clipping happens before the env sees the action (vecenv_wrapper.py:173-174)
        assert float(base.action_manager.action.abs().max()) <= 1.5 + 1e-6
    assert saw_timeout
    legacy = RslRlVecEnvWrapper(env.unwrapped, legacy_api=True)
    o, ex = legacy.get_observations()
    assert o.shape == (64, 48) and "observations" in ex
    log = env.episode_log()
    assert any(k.startswith("Episode_Reward/") for k in log) and any(k.startswith("Episode_Termination/") for k in log)


TRAIN_CFG = {
    "num_steps_per_env": 8, "save_interval": 50, "obs_groups": {"policy": ["policy"]},
    "policy": {"class_name": "ActorCritic", "init_noise_std": 1.0, "actor_hidden_dims": [64, 32],
               "critic_hidden_dims": [64, 32], "activation": "elu"},
    "algorithm": {"class_name": "PPO", "value_loss_coef": 1.0, "use_clipped_value_loss": True, "clip_param": 0.2,
                  "entropy_coef": 0.005, "num_learning_epochs": 2, "num_mini_batches": 2, "learning_rate": 1e-3,
                  "schedule": "adaptive", "gamma": 0.99, "lam": 0.95, "desired_kl": 0.01, "max_grad_norm": 1.0},
}


@pytest.mark.parametrize("task", ["Isaac-Cartpole-v0", "Isaac-Velocity-Rough-Anymal-C-v0"])
def test_runner_learn_save_load(task, tmp_path, monkeypatch):
    import copy
    from isaaclab_b200.rsl_rl import OnPolicyRunner, RslRlVecEnvWrapper
    monkeypatch.delenv("WORLD_SIZE", raising=False)
    env = RslRlVecEnvWrapper(_make_env(task, 128))
    runner = OnPolicyRunner(env, copy.deepcopy(TRAIN_CFG), log_dir=str(tmp_path), device="cuda")
    before = runner.alg.policy.flat_param.clone()
    runner.learn(3, init_at_random_ep_len=True)
    after = runner.alg.policy.flat_param
    assert torch.isfinite(after).all() and not torch.equal(before, after)
    assert runner.current_learning_iteration == 2 and runner.tot_timesteps == 3 * 8 * 128
    for k in ("Loss/value_function", "Loss/surrogate", "Loss/entropy", "Loss/learning_rate", "Policy/mean_noise_std",
              "Perf/total_fps", "Perf/collection time", "Perf/learning_time"):
        assert k in runner.last_log, k
    assert os.path.exists(tmp_path / "model_0.pt") and os.path.exists(tmp_path / "model_2.pt")
    ck = torch.load(tmp_path / "model_2.pt", weights_only=False)
    assert set(ck) == {"model_state_dict", "optimizer_state_dict", "iter", "infos"} and ck["iter"] == 2
    assert {"std", "actor.0.weight", "critic.0.weight"} <= set(ck["model_state_dict"])
    assert set(ck["optimizer_state_dict"]) == {"state", "param_groups"}

    env2 = RslRlVecEnvWrapper(_make_env(task, 128))
    r2 = OnPolicyRunner(env2, copy.deepcopy(TRAIN_CFG), log_dir=None, device="cuda")
    r2.load(str(tmp_path / "model_2.pt"))
    assert torch.equal(r2.alg.policy.flat_param, after) and r2.current_learning_iteration == 2
    assert torch.equal(r2.alg.exp_avg, runner.alg.exp_avg) and r2.alg.adam_step == runner.alg.adam_step
    policy = r2.get_inference_policy(device="cuda")
    obs = env2.get_observations()
    a1, a2 = policy(obs), runner.alg.policy.act_inference(obs)
    assert a1.shape == (128, env2.num_actions) and torch.equal(a1, a2)
    r2.learn(1)  # resumes without a log dir: no files, no log read
    assert r2.current_learning_iteration == 2
