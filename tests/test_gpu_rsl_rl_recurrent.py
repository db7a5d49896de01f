"""GPU parity of the recurrent policy path: `ActorCriticRecurrent` + `Memory` (LSTM / GRU, 1-2 layers)
forward in rollout mode, back-propagation through time in the PPO update, against fixtures recorded from the UNMODIFIED
rsl_rl (tests/golden/gen_rsl_rl.py --recurrent) and against torch.nn.LSTM / GRU autograd at a larger shape."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _mk(fx, dev="cuda"):
    from isaaclab_b200.rsl_rl import PPO, ActorCriticRecurrent, TensorDict
    c = fx["cfg"]
    obs0 = TensorDict({"policy": torch.zeros(c["N"], c["obs_dim"], device=dev)}, batch_size=[c["N"]])
    pol = ActorCriticRecurrent(obs0, {"policy": ["policy"], "critic": ["policy"]}, c["act_dim"], actor_hidden_dims=c["hidden"],
                               critic_hidden_dims=c["hidden"], device=dev, actor_obs_normalization=c["normalization"],
                               critic_obs_normalization=c["normalization"], **c["rnn"])
    assert list(pol.state_dict().keys()) == list(fx["state_dict_before"].keys())   # same names, same order as the reference
    pol.load_state_dict(fx["state_dict_before"])
    alg = PPO(pol, num_learning_epochs=c["epochs"], num_mini_batches=c["mini_batches"], clip_param=c["clip_param"],
              gamma=c["gamma"], lam=c["lam"], value_loss_coef=c["value_loss_coef"], entropy_coef=c["entropy_coef"],
              learning_rate=c["learning_rate"], max_grad_norm=c["max_grad_norm"], desired_kl=c["desired_kl"], device=dev)
    alg.init_storage("rl", c["N"], c["T"], obs0, [c["act_dim"]])
    return pol, alg, TensorDict


@pytest.mark.parametrize("name", ["lstm", "gru2", "lstm2"])
def test_recurrent_rollout_and_update_vs_reference_fixture(golden_dir, name):
    fx = torch.load(os.path.join(golden_dir, f"rsl_rl_{name}.pt"), weights_only=False)
    c = fx["cfg"]
    pol, alg, TensorDict = _mk(fx)
    dev = "cuda"
    for t, s in enumerate(fx["steps"]):
        obs = TensorDict({"policy": s["obs"].to(dev)}, batch_size=[c["N"]])
        eps = ((s["actions"] - s["mu"]) / s["sigma"]).to(dev)
        a = alg.act(obs, eps=eps)
        # rollout mode: one memory step from the carried hidden states, zeroed for finished envs (memory.py:36-53)
        torch.testing.assert_close(pol.action_mean.cpu(), s["mu"], rtol=1e-5, atol=3e-6)
        torch.testing.assert_close(alg.transition.values.cpu(), s["values"], rtol=1e-5, atol=3e-6)
        torch.testing.assert_close(a.cpu(), s["actions"], rtol=1e-5, atol=3e-6)
        extras = {} if s["time_outs"] is None else {"time_outs": s["time_outs"].to(dev)}
        nxt = fx["steps"][t + 1]["obs"] if t + 1 < len(fx["steps"]) else fx["last_obs"]
        alg.process_env_step(TensorDict({"policy": nxt.to(dev)}, batch_size=[c["N"]]), s["rewards"].to(dev),
                             s["dones"].to(dev), extras)
    st = alg.storage
    for mine, ref in ((st.saved_hidden_states_a, fx["storage"]["saved_hidden_states_a"]),
                      (st.saved_hidden_states_c, fx["storage"]["saved_hidden_states_c"])):
        assert len(mine) == len(ref)
        for a_, b_ in zip(mine, ref):
            torch.testing.assert_close(a_.cpu(), b_, rtol=1e-5, atol=3e-6)
    # the critic's evaluate() inside compute_returns advances its memory once more, as in the reference
    last_obs = TensorDict({"policy": fx["last_obs"].to(dev)}, batch_size=[c["N"]])
    torch.testing.assert_close(pol._evaluate(last_obs).cpu(), fx["last_values"], rtol=1e-5, atol=3e-6)
    # exact reference rollout from here on, so the update parity is not blurred by forward rounding
    for k in ("actions", "rewards", "values", "actions_log_prob", "mu", "sigma"):
        getattr(st, k).copy_(fx["storage"][k])
    for mine, ref in ((st.saved_hidden_states_a, fx["storage"]["saved_hidden_states_a"]),
                      (st.saved_hidden_states_c, fx["storage"]["saved_hidden_states_c"])):
        for a_, b_ in zip(mine, ref):
            a_.copy_(b_)
    st.compute_returns(fx["last_values"].to(dev), c["gamma"], c["lam"])
    torch.testing.assert_close(st.advantages.cpu(), fx["storage"]["advantages"], rtol=1e-5, atol=2e-6)
    loss = alg.update()
    for k, v in fx["state_dict_after"].items():
        got = pol.state_dict()[k].cpu()
        if not v.dtype.is_floating_point:
            assert torch.equal(got, v), k
            continue
        # same bounds as the feed-forward fixture test: Adam's sign-like step where |grad| ~ 1e-8
        torch.testing.assert_close(got, v, rtol=2e-4, atol=5e-4, msg=lambda m, k=k: f"{k}: {m}")
        tight = ((got - v).abs() <= 5e-6 + 2e-4 * v.abs()).float().mean().item()
        assert tight >= 0.99, f"{k}: only {tight:.4f} of the elements within rtol 2e-4 / atol 5e-6"
    for k in ("value_function", "surrogate", "entropy"):
        assert loss[k] == pytest.approx(fx["loss_dict"][k], rel=1e-4, abs=1e-6)
    assert alg.learning_rate == pytest.approx(float(fx["lr_per_update"][-1]), rel=1e-6)


@pytest.mark.parametrize("kind,layers", [("lstm", 1), ("gru", 1), ("lstm", 2), ("gru", 3)])
@pytest.mark.parametrize("precision", [0, 2])
def test_memory_bptt_vs_torch_autograd(kind, layers, precision):
    """Batch-mode forward + backward of `Memory` against torch.nn.LSTM / GRU autograd in fp64 on padded trajectories with
    ragged lengths, at a shape wide enough for the tcgen05 GEMMs (precision 2 = 3xTF32)."""
    from isaaclab_b200.rsl_rl import Memory, unpad_trajectories
    T, Btraj, I, H = 12, 96, 48, 128
    g = torch.Generator().manual_seed(5 + layers)
    mem = Memory(I, type=kind, num_layers=layers, hidden_size=H, device="cuda", precision=precision)
    for p_ in mem.parameters():
        p_.grad = torch.zeros_like(p_)
    ref = (torch.nn.LSTM if kind == "lstm" else torch.nn.GRU)(I, H, layers).double()
    ref.load_state_dict({k: v.detach().cpu().double() for k, v in mem.rnn.state_dict().items()})
    x = torch.randn(T, Btraj, I, generator=g)
    lengths = torch.randint(1, T + 1, (Btraj,), generator=g)
    lengths[0] = T
    # trajectories of one env are consecutive and fill T steps: build lengths that tile into rollouts of length T
    lens = []
    while len(lens) < Btraj:
        left = T
        while left > 0 and len(lens) < Btraj:
            n = int(torch.randint(1, left + 1, (1,), generator=g))
            lens.append(n)
            left -= n
        if left > 0:
            lens[-1] += left
    lengths = torch.tensor(lens)
    assert int(lengths.sum()) % T == 0
    masks = lengths.unsqueeze(0) > torch.arange(T).unsqueeze(1)
    x = x * masks.unsqueeze(-1)
    h0 = torch.randn(layers, Btraj, H, generator=g) * 0.5
    c0 = torch.randn(layers, Btraj, H, generator=g) * 0.5
    hid = [h0.cuda(), c0.cuda()] if kind == "lstm" else h0.cuda()
    out = mem(x.cuda(), masks.cuda(), hid)
    out_ref, _ = ref(x.double(), (h0.double(), c0.double()) if kind == "lstm" else h0.double())
    out_ref_u = unpad_trajectories(out_ref, masks)
    tol = dict(rtol=2e-5, atol=2e-5) if precision == 0 else dict(rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(out.cpu().double(), out_ref_u.detach(), **tol)
    d_out = torch.randn(out.shape, generator=g) / out.numel() ** 0.5
    out_ref_u.backward(d_out.double())
    mem.backward(d_out.cuda())
    for (k, p_), (_, q) in zip(mem.rnn.named_parameters(), ref.named_parameters()):
        torch.testing.assert_close(p_.grad.cpu().double(), q.grad, rtol=1e-3, atol=2e-5 if precision == 0 else 1e-4,
                                   msg=lambda m, k=k: f"{k}: {m}")
