"""Generate tests/golden/rsl_rl_*.pt by running the UNMODIFIED reference rsl_rl 3.0.0 on CPU.

Run in the authoring container only (`python tests/golden/gen_rsl_rl.py`); /root/reference must exist.
The fixture pins RolloutStorage.compute_returns, mini_batch_generator and PPO.update (reference:
source/rsl_rl/rsl_rl/storage/rollout_storage.py:127-203, algorithms/ppo.py:129-422).
"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness  # noqa: E402

ref_harness.install()
from rsl_rl.algorithms import PPO  # noqa: E402
from rsl_rl.modules import ActorCritic, ActorCriticRecurrent  # noqa: E402

TensorDict = ref_harness.TensorDict


def run(name, N, T, obs_dim, act_dim, hidden, seed, entropy_coef, epochs=2, mini_batches=4, timeouts=True,
        normalization=False, rnn=None, p_done=0.05):
    """`rnn`: dict(rnn_type, rnn_hidden_dim, rnn_num_layers) -> ActorCriticRecurrent (actor_critic_recurrent.py) and the
    recurrent minibatch generator (rollout_storage.py:206-260)."""
    torch.manual_seed(seed)
    torch.set_num_threads(1)
    obs0 = TensorDict({"policy": torch.randn(N, obs_dim)}, batch_size=[N])
    if rnn is not None:
        policy = ActorCriticRecurrent(obs0, {"policy": ["policy"], "critic": ["policy"]}, act_dim, actor_hidden_dims=hidden,
                                      critic_hidden_dims=hidden, activation="elu", init_noise_std=1.0,
                                      actor_obs_normalization=normalization, critic_obs_normalization=normalization, **rnn)
    else:
        policy = ActorCritic(obs0, {"policy": ["policy"], "critic": ["policy"]}, act_dim, actor_hidden_dims=hidden,
                             critic_hidden_dims=hidden, activation="elu", init_noise_std=1.0,
                             actor_obs_normalization=normalization, critic_obs_normalization=normalization)
    alg = PPO(policy, num_learning_epochs=epochs, num_mini_batches=mini_batches, clip_param=0.2, gamma=0.99, lam=0.95,
              value_loss_coef=1.0, entropy_coef=entropy_coef, learning_rate=1e-3, max_grad_norm=1.0,
              use_clipped_value_loss=True, schedule="adaptive", desired_kl=0.01, device="cpu")
    alg.init_storage("rl", N, T, obs0, [act_dim])
    out = {"cfg": dict(N=N, T=T, obs_dim=obs_dim, act_dim=act_dim, hidden=list(hidden), entropy_coef=entropy_coef,
                       epochs=epochs, mini_batches=mini_batches, gamma=0.99, lam=0.95, clip_param=0.2,
                       value_loss_coef=1.0, learning_rate=1e-3, max_grad_norm=1.0, desired_kl=0.01,
                       normalization=normalization, rnn=rnn)}
    out["state_dict_before"] = {k: v.clone() for k, v in policy.state_dict().items()}
    g = torch.Generator().manual_seed(seed + 1)
    col_scale = torch.ones(obs_dim) if not normalization else 0.5 + 3.0 * torch.rand(obs_dim, generator=g)
    col_shift = torch.zeros(obs_dim) if not normalization else 2.0 * torch.randn(obs_dim, generator=g)
    obs = obs0
    step_inputs = []
    for t in range(T):
        with torch.inference_mode():
            actions = alg.act(obs)
        rewards = torch.randn(N, generator=g) * 0.05
        dones = (torch.rand(N, generator=g) < p_done).long()
        time_outs = ((torch.rand(N, generator=g) < 0.3) & (dones > 0)) if timeouts else None
        # observations with per-column offset / scale so the empirical normalisation has something to learn
        next_obs = TensorDict({"policy": torch.randn(N, obs_dim, generator=g) * col_scale + col_shift}, batch_size=[N])
        step_inputs.append(dict(obs=obs["policy"].clone(), actions=actions.clone(), rewards=rewards.clone(),
                                dones=dones.clone(), time_outs=None if time_outs is None else time_outs.clone(),
                                values=alg.transition.values.clone(), log_prob=alg.transition.actions_log_prob.clone(),
                                mu=alg.transition.action_mean.clone(), sigma=alg.transition.action_sigma.clone()))
        extras = {"time_outs": time_outs} if timeouts else {}
        with torch.inference_mode():
            alg.process_env_step(next_obs, rewards, dones, extras)
        obs = next_obs
    out["steps"] = step_inputs
    out["last_obs"] = obs["policy"].clone()
    st = alg.storage
    seen = {}
    orig_cr = st.compute_returns

    def compute_returns(last_values, *a, **k):   # a recurrent critic's evaluate() advances its memory: record THE call's values
        seen["last_values"] = last_values.detach().clone()
        return orig_cr(last_values, *a, **k)

    st.compute_returns = compute_returns
    with torch.inference_mode():
        alg.compute_returns(obs)
    st.compute_returns = orig_cr
    out["storage"] = dict(observations=st.observations["policy"].clone(), actions=st.actions.clone(),
                          rewards=st.rewards.clone(), dones=st.dones.clone(), values=st.values.clone(),
                          actions_log_prob=st.actions_log_prob.clone(), mu=st.mu.clone(), sigma=st.sigma.clone(),
                          returns=st.returns.clone(), advantages=st.advantages.clone())
    out["last_values"] = seen["last_values"]
    if rnn is not None:   # [T, num_layers, N, H] per state tensor: (h, c) for an LSTM, (h,) for a GRU
        out["storage"]["saved_hidden_states_a"] = [h.clone() for h in st.saved_hidden_states_a]
        out["storage"]["saved_hidden_states_c"] = [h.clone() for h in st.saved_hidden_states_c]
    if rnn is not None:   # what the reference's recurrent generator yields (first epoch), for the host-side test
        def cl(v):
            return [x.clone() for x in v] if isinstance(v, (list, tuple)) else v.clone()
        out["recurrent_batches"] = [dict(obs=mb[0]["policy"].clone(), actions=mb[1].clone(), values=mb[2].clone(),
                                         advantages=mb[3].clone(), returns=mb[4].clone(), log_prob=mb[5].clone(),
                                         mu=mb[6].clone(), sigma=mb[7].clone(), hid_a=cl(mb[8][0]), hid_c=cl(mb[8][1]),
                                         masks=mb[9].clone())
                                    for mb in st.recurrent_mini_batch_generator(mini_batches, 1)]
    # record the permutation the generator draws (rollout_storage.py:165)
    recorded = {}
    orig_randperm = torch.randperm

    def randperm(*a, **k):
        r = orig_randperm(*a, **k)
        recorded["indices"] = r.clone()
        return r

    torch.randperm = randperm
    lrs = []
    orig_step = alg.optimizer.step

    def step(*a, **k):
        lrs.append(alg.learning_rate)
        return orig_step(*a, **k)

    alg.optimizer.step = step
    try:
        loss = alg.update()
    finally:
        torch.randperm = orig_randperm
    out["indices"] = recorded.get("indices")   # the recurrent generator draws no permutation
    out["lr_per_update"] = torch.tensor(lrs, dtype=torch.float64)
    out["loss_dict"] = {k: float(v) for k, v in loss.items()}
    out["state_dict_after"] = {k: v.clone() for k, v in policy.state_dict().items()}
    opt = alg.optimizer.state_dict()["state"]
    names = [n for n, _ in policy.named_parameters()]
    out["adam_after"] = {names[i]: dict(exp_avg=s["exp_avg"].clone(), exp_avg_sq=s["exp_avg_sq"].clone(),
                                        step=float(s["step"])) for i, s in opt.items()}
    path = os.path.join(HERE, f"rsl_rl_{name}.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB; loss", out["loss_dict"], "lr", lrs[-1])


def main_recurrent():
    run("lstm", N=24, T=8, obs_dim=14, act_dim=4, hidden=[32, 32], seed=31, entropy_coef=0.005, epochs=2, mini_batches=3,
        rnn=dict(rnn_type="lstm", rnn_hidden_dim=24, rnn_num_layers=1), p_done=0.12)
    run("gru2", N=16, T=6, obs_dim=10, act_dim=3, hidden=[16], seed=37, entropy_coef=0.01, epochs=1, mini_batches=2,
        rnn=dict(rnn_type="gru", rnn_hidden_dim=20, rnn_num_layers=2), p_done=0.15, normalization=True)
    run("lstm2", N=12, T=5, obs_dim=9, act_dim=2, hidden=[16, 16], seed=41, entropy_coef=0.0, epochs=1, mini_batches=2,
        rnn=dict(rnn_type="lstm", rnn_hidden_dim=12, rnn_num_layers=2), p_done=0.2, timeouts=False)


if __name__ == "__main__":
    if "--recurrent" in sys.argv:
        main_recurrent()
        sys.exit(0)
    run("small", N=96, T=8, obs_dim=48, act_dim=12, hidden=[64, 64, 64], seed=7, entropy_coef=0.005)
    run("ragged", N=50, T=5, obs_dim=37, act_dim=3, hidden=[32, 16], seed=11, entropy_coef=0.01, epochs=1,
        mini_batches=3, timeouts=False)
    run("norm", N=64, T=6, obs_dim=20, act_dim=4, hidden=[32, 32], seed=23, entropy_coef=0.005, epochs=2, mini_batches=2,
        normalization=True)
