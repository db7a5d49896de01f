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


def test_rollout_stores_the_observation_act_saw(monkeypatch):
    """the fused env returns ONE persistent observation buffer and overwrites it in `step`, so the
    rollout slab must hold obs_t -- the observation `PPO.act` saw -- next to a_t / log_prob_t / mu_t / V_t, not obs_{t+1}.
    Checked on the runner's own loop: storage.observations[t] == the tensor passed to act() at step t, and the policy
    re-evaluated on the stored rows reproduces the stored mu / values / log-prob (importance ratio 1 before any update)."""
    import copy
    from isaaclab_b200.rsl_rl import OnPolicyRunner, RslRlVecEnvWrapper
    from isaaclab_b200.rsl_rl.tensordict import TensorDict
    monkeypatch.delenv("WORLD_SIZE", raising=False)
    env = RslRlVecEnvWrapper(_make_env("Isaac-Velocity-Flat-Anymal-C-v0", 96))
    runner = OnPolicyRunner(env, copy.deepcopy(TRAIN_CFG), log_dir=None, device="cuda")
    alg, T = runner.alg, TRAIN_CFG["num_steps_per_env"]
    obs = env.get_observations()
    seen = []
    for t in range(T):
        seen.append(obs["policy"].clone())
        actions = alg.act(obs)
        obs, rew, dones, extras = env.step(actions)
        assert not torch.equal(obs["policy"], seen[-1]), "env.step did not change the observation buffer"
        alg.process_env_step(obs, rew, dones, extras)
    st = alg.storage
    for t in range(T):
        assert torch.equal(st.observations["policy"][t], seen[t]), f"storage row {t} is not the observation act() saw"
    pol = alg.policy
    flat = TensorDict({"policy": st.observations["policy"].flatten(0, 1).contiguous()}, batch_size=[T * 96])
    mu = pol.act_inference(flat)
    val = pol.evaluate(flat)
    torch.testing.assert_close(mu, st.mu.flatten(0, 1), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(val, st.values.flatten(0, 1), rtol=1e-5, atol=1e-6)
    sigma = pol.std.detach()
    lp = (-((st.actions.flatten(0, 1) - mu) ** 2) / (2 * sigma ** 2) - torch.log(sigma) - 0.9189385332046727).sum(-1)
    ratio = torch.exp(lp - st.actions_log_prob.flatten(0, 1).squeeze(-1))
    torch.testing.assert_close(ratio, torch.ones_like(ratio), rtol=0, atol=2e-5)


def test_policy_entry_points_return_fresh_tensors():
    """act_inference / evaluate results must survive the next call with the same batch size."""
    from isaaclab_b200.rsl_rl import ActorCritic
    from isaaclab_b200.rsl_rl.tensordict import TensorDict
    torch.manual_seed(0)
    o1 = TensorDict({"policy": torch.randn(32, 10, device="cuda")}, batch_size=[32])
    o2 = TensorDict({"policy": torch.randn(32, 10, device="cuda")}, batch_size=[32])
    pol = ActorCritic(o1, {"policy": ["policy"], "critic": ["policy"]}, 3, actor_hidden_dims=[16], critic_hidden_dims=[16],
                      device="cuda")
    a1, v1 = pol.act_inference(o1), pol.evaluate(o1)
    a1c, v1c = a1.clone(), v1.clone()
    pol.act_inference(o2), pol.evaluate(o2)
    assert torch.equal(a1, a1c) and torch.equal(v1, v1c)


def test_asymmetric_critic_minibatches_carry_every_group():
    """with obs_groups['critic'] = ['critic'] the minibatch TensorDict must hold both groups."""
    from isaaclab_b200.rsl_rl import PPO, ActorCritic
    from isaaclab_b200.rsl_rl.tensordict import TensorDict
    N, T = 64, 4
    g = torch.Generator(device="cuda").manual_seed(1)
    mk = lambda: TensorDict({"policy": torch.randn(N, 9, device="cuda", generator=g),   # noqa: E731
                             "critic": torch.randn(N, 14, device="cuda", generator=g)}, batch_size=[N])
    obs = mk()
    pol = ActorCritic(obs, {"policy": ["policy"], "critic": ["critic"]}, 3, actor_hidden_dims=[16], critic_hidden_dims=[16],
                      device="cuda")
    alg = PPO(pol, num_learning_epochs=1, num_mini_batches=2, device="cuda")
    alg.init_storage("rl", N, T, obs, [3])
    rows = []
    for _ in range(T):
        rows.append({k: obs[k].clone() for k in ("policy", "critic")})
        alg.act(obs)
        obs = mk()
        alg.process_env_step(obs, torch.randn(N, device="cuda", generator=g), torch.zeros(N, dtype=torch.long, device="cuda"), {})
    for t in range(T):
        for k in ("policy", "critic"):
            assert torch.equal(alg.storage.observations[k][t], rows[t][k])
    alg.compute_returns(obs)
    idx = torch.randperm(N * T, device="cuda")
    for batch in alg.storage.mini_batch_generator(2, 1, indices=idx):
        ob = batch[0]
        assert set(ob.keys()) == {"policy", "critic"} and ob["critic"].shape == (N * T // 2, 14)
        break
    first = idx[: N * T // 2]
    torch.testing.assert_close(ob["critic"], alg.storage.observations["critic"].flatten(0, 1)[first])
    loss = alg.update(indices=idx)
    assert all(torch.isfinite(torch.tensor(v)) for v in loss.values())


@pytest.mark.parametrize("rnn_type,layers", [("lstm", 1), ("gru", 2)])
def test_runner_learns_with_recurrent_policy(rnn_type, layers, tmp_path, monkeypatch):
    """`class_name: "ActorCriticRecurrent"` through the runner's own loop -- rollout with carried and
    reset hidden states, recurrent minibatches, BPTT update, checkpoint round trip (`memory_a.rnn.*` keys)."""
    import copy
    from isaaclab_b200.rsl_rl import OnPolicyRunner, RslRlVecEnvWrapper
    monkeypatch.delenv("WORLD_SIZE", raising=False)
    cfg = copy.deepcopy(TRAIN_CFG)
    cfg["policy"].update(class_name="ActorCriticRecurrent", rnn_type=rnn_type, rnn_hidden_dim=64, rnn_num_layers=layers)
    env = RslRlVecEnvWrapper(_make_env("Isaac-Velocity-Flat-Anymal-C-v0", 128))
    runner = OnPolicyRunner(env, copy.deepcopy(cfg), log_dir=str(tmp_path), device="cuda")
    pol = runner.alg.policy
    assert pol.is_recurrent and f"memory_a.rnn.weight_hh_l{layers - 1}" in pol.state_dict()
    before = pol.flat_param.clone()
    runner.learn(3, init_at_random_ep_len=True)
    assert torch.isfinite(pol.flat_param).all() and not torch.equal(before, pol.flat_param)
    rnn_slice = pol.memory_a.rnn.weight_hh_l0
    assert not torch.equal(rnn_slice, torch.load(tmp_path / "model_0.pt", weights_only=False)["model_state_dict"]
                           ["memory_a.rnn.weight_hh_l0"].to(rnn_slice.device))   # the RNN weights were trained too
    h = pol.get_hidden_states()[0]
    h = h[0] if isinstance(h, tuple) else h
    assert h.shape == (layers, 128, 64) and torch.isfinite(h).all()
    r2 = OnPolicyRunner(RslRlVecEnvWrapper(_make_env("Isaac-Velocity-Flat-Anymal-C-v0", 128)), copy.deepcopy(cfg), log_dir=None,
                        device="cuda")
    r2.load(str(tmp_path / "model_2.pt"))
    assert torch.equal(r2.alg.policy.flat_param, pol.flat_param)
