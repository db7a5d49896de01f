"""Host-side mirror of the reference interface: name resolution against golden answers of the unmodified reference
(tests/golden/gen_host.py), env-cfg compilation of the five BASELINE tasks, error behaviour."""
import json
import os

import pytest

from isaaclab_b200.utils import string as S
from tests.mdp_common import make_spec

FX = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "host_logic.json")))


def test_resolve_matching_names_matches_reference():
    for c in FX["names"]:
        idx, names = S.resolve_matching_names(c["keys"], c["strings"], preserve_order=c["preserve_order"])
        assert (idx, names) == (c["idx"], c["names"]), c["keys"]


def test_resolve_matching_names_errors_like_reference():
    for c in FX["errors"]:
        if c["raises"]:
            with pytest.raises(ValueError):
                S.resolve_matching_names(c["keys"], c["strings"])
        else:
            S.resolve_matching_names(c["keys"], c["strings"])


def test_resolve_matching_names_values_matches_reference():
    for c in FX["values"]:
        idx, names, vals = S.resolve_matching_names_values(c["data"], c["strings"], preserve_order=c["preserve_order"])
        assert (idx, names, vals) == (c["idx"], c["names"], c["values"]), c["data"]


# observation widths / term counts of the reference task cfgs (velocity_env_cfg.py:127-158, rough_env_cfg.py of each
# robot, cartpole_env_cfg.py:80-100); the step-level behaviour of the same cfgs is pinned by test_oracle_mdp_golden.py
EXPECT = {
    "Isaac-Cartpole-v0": dict(J=2, A=1, obs=4, rewards=5, terminations=2, R=0),
    "Isaac-Velocity-Flat-Anymal-C-v0": dict(J=12, A=12, obs=48, rewards=11, terminations=2, R=0),
    "Isaac-Velocity-Rough-Anymal-C-v0": dict(J=12, A=12, obs=235, rewards=11, terminations=2, R=187),
    "Isaac-Velocity-Rough-Unitree-Go1-v0": dict(J=12, A=12, obs=235, rewards=10, terminations=2, R=187),  # undesired_contacts = None
    "Isaac-Velocity-Rough-H1-v0": dict(J=19, A=19, obs=256, R=187),
}


@pytest.mark.parametrize("task", list(EXPECT))
def test_compile_env_cfg(task):
    from isaaclab_b200.envs.term_dims import term_dim
    spec = make_spec(task, 16)
    want = EXPECT[task]
    assert spec["J"] == want["J"] and spec["A"] == want["A"] and spec.get("R", 0) == want["R"]
    assert sum(term_dim(spec, t) for t in spec["observations"]["terms"]) == want["obs"]
    if "rewards" in want:
        assert len(spec["rewards"]) == want["rewards"]
        assert len(spec["terminations"]) == want["terminations"]
    import math
    assert spec["max_episode_length"] == math.ceil(spec["max_episode_length_s"] / spec["step_dt"])  # manager_based_rl_env.py:97


def test_invalid_action_shape_raises_like_action_manager():
    """ActionManager.process_action raises ValueError on a wrong action width (action_manager.py:325-327); the
    mirror checks before touching the device."""
    from isaaclab_b200.envs.fused_env import FusedEnv
    import torch
    env = FusedEnv.__new__(FusedEnv)
    env.N, env.A = 4, 12
    with pytest.raises(ValueError, match="Invalid action shape"):
        FusedEnv.process_action(env, torch.zeros(4, 11))
