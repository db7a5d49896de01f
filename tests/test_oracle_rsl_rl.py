"""The rsl_rl oracle restatement (oracle/rsl_rl.py) against fixtures produced by the UNMODIFIED reference
(tests/golden/gen_rsl_rl.py).  CPU only."""
import os

import pytest
import torch

from oracle import rsl_rl as O


@pytest.fixture(params=["small", "ragged"])
def fx(request, golden_dir):
    return torch.load(os.path.join(golden_dir, f"rsl_rl_{request.param}.pt"), weights_only=False)


def test_gae_matches_reference(fx):
    st, c = fx["storage"], fx["cfg"]
    ret, adv = O.compute_returns(st["rewards"], st["values"], st["dones"], fx["last_values"], c["gamma"], c["lam"])
    torch.testing.assert_close(ret, st["returns"], rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(adv, st["advantages"], rtol=1e-5, atol=1e-6)


def test_timeout_bootstrap_and_storage(fx):
    c = fx["cfg"]
    for t, s in enumerate(fx["steps"]):
        r = s["rewards"] if s["time_outs"] is None else O.bootstrap_rewards(s["rewards"], s["values"], s["time_outs"], c["gamma"])
        torch.testing.assert_close(r.view(-1, 1), fx["storage"]["rewards"][t], rtol=0, atol=0)


def test_act_matches_reference(fx):
    p = fx["state_dict_before"]
    s = fx["steps"][0]
    mu = O.mlp_forward(p, "actor", s["obs"], O.n_layers_of(p, "actor"))
    torch.testing.assert_close(mu, s["mu"], rtol=1e-5, atol=1e-6)
    v = O.mlp_forward(p, "critic", s["obs"], O.n_layers_of(p, "critic"))
    torch.testing.assert_close(v, s["values"], rtol=1e-5, atol=1e-6)
    lp = O.normal_log_prob(s["actions"], mu, p["std"].expand_as(mu))
    torch.testing.assert_close(lp, s["log_prob"], rtol=1e-5, atol=1e-5)


def test_ppo_update_matches_reference(fx):
    c = fx["cfg"]
    params = {k: v.clone() for k, v in fx["state_dict_before"].items()}
    loss, lr, lrs, adam = O.ppo_update(params, fx["storage"], fx["indices"], c, c["learning_rate"])
    assert lrs == pytest.approx(fx["lr_per_update"].tolist(), rel=1e-12)
    for k, v in fx["state_dict_after"].items():
        torch.testing.assert_close(params[k], v, rtol=1e-4, atol=2e-6, msg=lambda m, k=k: f"{k}: {m}")
    for k in ("value_function", "surrogate", "entropy"):
        assert loss[k] == pytest.approx(fx["loss_dict"][k], rel=1e-5, abs=1e-7)
    for k, s in fx["adam_after"].items():
        torch.testing.assert_close(adam.m[k], s["exp_avg"], rtol=1e-4, atol=1e-7)


def test_empirical_normalization_matches_reference(golden_dir):
    """oracle EmpiricalNormalization against the statistics the reference accumulated over the 'norm' rollout, and its
    normalised forward against the reference actor's recorded means."""
    import os
    fx = torch.load(os.path.join(golden_dir, "rsl_rl_norm.pt"), weights_only=False)
    c = fx["cfg"]
    norm = O.EmpiricalNormalization(c["obs_dim"])
    seen = [s["obs"] for s in fx["steps"][1:]] + [fx["last_obs"]]   # process_env_step learns the NEXT observation
    for t, s in enumerate(fx["steps"]):
        # the reference's act(obs_t) normalises with the statistics learnt so far
        sd = fx["state_dict_before"]
        x = norm(s["obs"])
        for i in range(len(c["hidden"]) + 1):
            x = torch.nn.functional.linear(x, sd[f"actor.{2 * i}.weight"], sd[f"actor.{2 * i}.bias"])
            if i < len(c["hidden"]):
                x = torch.nn.functional.elu(x)
        torch.testing.assert_close(x, s["mu"], rtol=1e-5, atol=1e-6)
        norm.update(seen[t])
    after = fx["state_dict_after"]
    assert norm.count == int(after["actor_obs_normalizer.count"])
    torch.testing.assert_close(norm.mean, after["actor_obs_normalizer._mean"], rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(norm.var, after["actor_obs_normalizer._var"], rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(norm.std, after["actor_obs_normalizer._std"], rtol=1e-6, atol=1e-7)
