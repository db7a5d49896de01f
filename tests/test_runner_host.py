"""Host-side logic of the rsl_rl runner mirror (no GPU): observation-set resolution follows RSL/utils/utils.py:202-309,
multi-GPU environment parsing follows RSL/runners/on_policy_runner.py:353-395."""
import warnings

import pytest
import torch

from isaaclab_b200.rsl_rl.runner import OnPolicyRunner, resolve_obs_groups
from isaaclab_b200.rsl_rl.tensordict import TensorDict


def _obs(**groups):
    return TensorDict({k: torch.zeros(4, d) for k, d in groups.items()}, batch_size=[4])


def test_resolve_obs_groups_defaults_and_errors():
    obs = _obs(policy=5, critic=7, extra=2)
    out = resolve_obs_groups(obs, {"policy": ["policy", "extra"]}, ["critic"])
    assert out == {"policy": ["policy", "extra"], "critic": ["critic"]}  # same-named group wins
    out = resolve_obs_groups(_obs(policy=5), {"policy": ["policy"]}, ["critic"])
    assert out["critic"] == ["policy"] and out["critic"] is not out["policy"]  # falls back to a COPY of the policy set
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        assert resolve_obs_groups(_obs(policy=5), {}, ["critic"])["policy"] == ["policy"]
    with pytest.raises(ValueError, match="must contain the 'policy' key"):
        resolve_obs_groups(_obs(state=5), {}, ["critic"])
    with pytest.raises(ValueError, match="can not be an empty list"):
        resolve_obs_groups(obs, {"policy": []}, ["critic"])
    with pytest.raises(ValueError, match="not found in the observations"):
        resolve_obs_groups(obs, {"policy": ["nope"]}, ["critic"])


def test_multi_gpu_env_validation(monkeypatch):
    r = object.__new__(OnPolicyRunner)
    r.device = "cuda:0"
    monkeypatch.delenv("WORLD_SIZE", raising=False)
    r._configure_multi_gpu()
    assert not r.is_distributed and r.multi_gpu_cfg is None and r.gpu_world_size == 1
    monkeypatch.setenv("WORLD_SIZE", "2")
    monkeypatch.setenv("LOCAL_RANK", "1")
    monkeypatch.setenv("RANK", "1")
    with pytest.raises(ValueError, match="does not match expected device for local rank"):
        r._configure_multi_gpu()
    r.device = "cuda:3"
    monkeypatch.setenv("LOCAL_RANK", "3")
    with pytest.raises(ValueError, match="greater than or equal to world size"):
        r._configure_multi_gpu()


def test_unknown_classes_fail_loudly():
    r = object.__new__(OnPolicyRunner)
    r.cfg = {"obs_groups": {"policy": ["policy"], "critic": ["policy"]}}
    r.device, r.multi_gpu_cfg = "cpu", None
    r.policy_cfg = {"class_name": "StudentTeacher"}
    r.alg_cfg = {"class_name": "PPO"}
    with pytest.raises(NotImplementedError, match="StudentTeacher"):
        r._construct_algorithm(_obs(policy=5))
    r.alg_cfg = {"class_name": "PPO", "rnd_cfg": {"weight": 1.0}}
    with pytest.raises(NotImplementedError, match="RND"):
        r._construct_algorithm(_obs(policy=5))
