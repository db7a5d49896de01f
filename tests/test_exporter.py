"""Policy exporter (RL/rsl_rl/exporter.py): the TorchScript artefact reproduces the policy's inference path on CPU, carries
its memory for recurrent policies and resets it."""
import torch
import torch.nn.functional as F

from isaaclab_b200.rsl_rl import ActorCritic, ActorCriticRecurrent, TensorDict
from isaaclab_b200.rsl_rl.exporter import export_policy_as_jit


def _mlp_ref(seq, x):
    lins = [m for m in seq if isinstance(m, torch.nn.Linear)]
    for i, l in enumerate(lins):
        x = F.linear(x, l.weight.detach(), l.bias.detach())
        if i + 1 < len(lins):
            x = F.elu(x)
    return x


def test_feedforward_policy_exports_with_normalizer(tmp_path):
    torch.manual_seed(0)
    obs0 = TensorDict({"policy": torch.zeros(4, 37)}, batch_size=[4])   # 37: first-layer weight lives in a K-padded block
    pol = ActorCritic(obs0, {"policy": ["policy"], "critic": ["policy"]}, 5, actor_hidden_dims=[16, 8], critic_hidden_dims=[8],
                      device="cpu", actor_obs_normalization=True)
    assert pol.actor[0].weight.stride(0) == 40 and pol.actor[0].weight.shape == (16, 37)
    pol.actor_obs_normalizer._mean.copy_(torch.randn(1, 37))
    pol.actor_obs_normalizer._std.copy_(torch.rand(1, 37) + 0.5)
    export_policy_as_jit(pol, pol.actor_obs_normalizer, str(tmp_path), "policy.pt")
    m = torch.jit.load(str(tmp_path / "policy.pt"))
    x = torch.randn(9, 37)
    n = pol.actor_obs_normalizer
    want = _mlp_ref(pol.actor, (x - n._mean) / (n._std + n.eps))
    torch.testing.assert_close(m(x), want, rtol=1e-6, atol=1e-6)
    m.reset()


def test_recurrent_policy_exports_its_memory(tmp_path):
    torch.manual_seed(1)
    obs0 = TensorDict({"policy": torch.zeros(2, 10)}, batch_size=[2])
    for kind in ("lstm", "gru"):
        pol = ActorCriticRecurrent(obs0, {"policy": ["policy"], "critic": ["policy"]}, 3, actor_hidden_dims=[8],
                                   critic_hidden_dims=[8], rnn_type=kind, rnn_hidden_dim=12, rnn_num_layers=2, device="cpu")
        export_policy_as_jit(pol, None, str(tmp_path), f"{kind}.pt")
        m = torch.jit.load(str(tmp_path / f"{kind}.pt"))
        ref = (torch.nn.LSTM if kind == "lstm" else torch.nn.GRU)(10, 12, 2)
        ref.load_state_dict({k: v.detach().clone() for k, v in pol.memory_a.rnn.state_dict().items()})
        xs = torch.randn(4, 1, 10)
        y, _ = ref(xs)                                    # 4 steps of one env from a zero state
        outs = [m(xs[t]) for t in range(4)]
        for t in range(4):
            torch.testing.assert_close(outs[t], _mlp_ref(pol.actor, y[t]), rtol=1e-5, atol=1e-6)
        m.reset()
        torch.testing.assert_close(m(xs[0]), outs[0], rtol=1e-6, atol=1e-7)   # state was cleared
