"""Contract 1 at construction time, without a GPU: `ManagerBase._resolve_common_term_cfg` semantics
(LAB/managers/manager_base.py:278-351) in `isaaclab_b200.envs.compile` -- "module:attr" strings, the signature check,
class terms, the reference's own mdp functions recognised as fused identities, and the Python-term (EXTERNAL) fallback."""
import pytest
import torch

from isaaclab_b200.envs import mdp
from isaaclab_b200.envs.compile import compile_env_cfg
from isaaclab_b200.managers import ManagerTermBase, ObservationTermCfg, RewardTermCfg, SceneEntityCfg, TerminationTermCfg
from isaaclab_b200.tasks import TASKS

TASK = "Isaac-Velocity-Flat-Anymal-C-v0"


def user_reward(env, scale: float, asset_cfg: SceneEntityCfg = SceneEntityCfg("robot")):
    return scale * torch.sum(torch.square(env.scene[asset_cfg.name].data.joint_vel[:, asset_cfg.joint_ids]), dim=1)


class CountingTerm(ManagerTermBase):
    def __call__(self, env, gain: float = 1.0):
        return torch.zeros(env.num_envs)


def _cfg():
    return TASKS[TASK]()


def test_user_function_becomes_an_external_term_and_scene_entities_are_resolved():
    cfg = _cfg()
    cfg.rewards.user = RewardTermCfg(func=user_reward, weight=-0.5,
                                     params={"scale": 2.0, "asset_cfg": SceneEntityCfg("robot", joint_names=[".*HFE"])})
    spec = compile_env_cfg(cfg, num_envs=4)
    t = [r for r in spec["rewards"] if r["name"] == "user"][0]
    assert t["op"] == "external" and t["weight"] == -0.5 and t["func"] is user_reward and not t["is_class"]
    assert t["params"]["asset_cfg"].joint_ids == [i for i, n in enumerate(spec["joint_names"]) if n.endswith("HFE")]
    fused = [r for r in spec["rewards"] if r["op"] != "external"]
    assert len(fused) == len(spec["rewards"]) - 1


def test_string_funcs_resolve_like_string_to_callable():
    cfg = _cfg()
    cfg.rewards.user = RewardTermCfg(func="tests.test_term_resolution:user_reward", weight=1.0, params={"scale": 1.0})
    cfg.terminations.t2 = TerminationTermCfg(func="isaaclab_b200.envs.mdp:bad_orientation", params={"limit_angle": 1.0})
    spec = compile_env_cfg(cfg, num_envs=4)
    assert [r for r in spec["rewards"] if r["name"] == "user"][0]["func"] is user_reward
    assert [t for t in spec["terminations"] if t["name"] == "t2"][0]["op"] == "bad_orientation"
    cfg = _cfg()
    cfg.rewards.bad = RewardTermCfg(func="no.such.module:fn", weight=1.0)
    with pytest.raises(ValueError, match="Could not resolve the input string"):
        compile_env_cfg(cfg, num_envs=4)


def test_signature_is_checked_against_params_with_the_reference_message():
    cfg = _cfg()
    cfg.rewards.user = RewardTermCfg(func=user_reward, weight=1.0, params={})            # `scale` missing
    with pytest.raises(ValueError, match=r"expects mandatory parameters: \['scale'\] and optional parameters: \['asset_cfg'\]"):
        compile_env_cfg(cfg, num_envs=4)
    cfg = _cfg()
    cfg.rewards.user = RewardTermCfg(func=user_reward, weight=1.0, params={"scale": 1.0, "typo": 3})
    with pytest.raises(ValueError, match="but received"):
        compile_env_cfg(cfg, num_envs=4)
    cfg = _cfg()
    cfg.rewards.user = RewardTermCfg(func=3.14, weight=1.0)
    with pytest.raises(AttributeError, match="is not callable"):
        compile_env_cfg(cfg, num_envs=4)
    cfg = _cfg()
    cfg.rewards.user = RewardTermCfg(func=user_reward, weight="heavy", params={"scale": 1.0})
    with pytest.raises(TypeError, match="is not of type float or int"):
        compile_env_cfg(cfg, num_envs=4)


def test_class_terms_must_derive_from_manager_term_base():
    cfg = _cfg()
    cfg.rewards.cls = RewardTermCfg(func=CountingTerm, weight=1.0, params={"gain": 2.0})
    spec = compile_env_cfg(cfg, num_envs=4)
    assert [r for r in spec["rewards"] if r["name"] == "cls"][0]["is_class"]

    class NotATerm:
        def __call__(self, env):
            return None

    cfg = _cfg()
    cfg.rewards.cls = RewardTermCfg(func=NotATerm, weight=1.0)
    with pytest.raises(TypeError, match="is not of type ManagerTermBase"):
        compile_env_cfg(cfg, num_envs=4)


def test_user_observation_term_width_is_deferred_to_env_construction():
    cfg = _cfg()
    cfg.observations.policy.extra = ObservationTermCfg(func=lambda env: env.scene["robot"].data.joint_pos[:, :3])
    spec = compile_env_cfg(cfg, num_envs=4)
    t = spec["observations"]["terms"][-1]
    assert t["op"] == "external" and t["dim"] is None


def test_events_and_curriculum_stay_on_the_fused_library():
    from isaaclab_b200.managers import EventTermCfg
    cfg = _cfg()
    cfg.events.mine = EventTermCfg(func=lambda env, env_ids: None, mode="reset")
    with pytest.raises(NotImplementedError, match="user event terms"):
        compile_env_cfg(cfg, num_envs=4)


def test_reference_mdp_functions_compile_to_the_same_opcodes():
    """A cfg that passes the REFERENCE's own function objects (isaaclab.envs.mdp.*, the velocity task's mdp package) lowers
    to the same term programs as one that passes isaaclab_b200.envs.mdp identities."""
    from tests.golden import ref_harness
    if not ref_harness.available():
        pytest.skip("reference sources not present")
    ref_harness.install()
    import importlib
    ref_mdp = importlib.import_module("isaaclab.envs.mdp")
    vel_mdp = importlib.import_module("isaaclab_tasks.manager_based.locomotion.velocity.mdp")
    want = compile_env_cfg(_cfg(), num_envs=4)
    cfg = _cfg()
    swapped = 0
    for container in (cfg.rewards, cfg.terminations, cfg.observations.policy):
        for name, t in vars(container).items():
            f = getattr(t, "func", None)
            if not hasattr(f, "_b200_op"):
                continue
            ref_f = getattr(vel_mdp, f._b200_op, None) or getattr(ref_mdp, f._b200_op, None)
            assert ref_f is not None and not hasattr(ref_f, "_b200_op"), f._b200_op
            t.func = ref_f
            swapped += 1
    assert swapped >= 15
    got = compile_env_cfg(cfg, num_envs=4)
    for key in ("rewards", "terminations"):
        assert [(t["name"], t["op"]) for t in got[key]] == [(t["name"], t["op"]) for t in want[key]]
        assert all(t["op"] != "external" for t in got[key])
    assert [t["op"] for t in got["observations"]["terms"]] == [t["op"] for t in want["observations"]["terms"]]
    assert mdp._REGISTRY  # the registry the recognition goes through


def test_history_and_modifiers_compile_like_the_reference():
    """observation_manager.py:411-433: the group's history_length overrides the terms'; a term with history occupies
    dim x history_length columns; function modifiers compile in order; class modifiers are refused with a reason."""
    from isaaclab_b200.envs.compile import compile_env_cfg
    from isaaclab_b200.envs.term_dims import term_dim, term_raw_dim
    from isaaclab_b200.tasks import TASKS
    from isaaclab_b200.utils import modifiers
    cfg = TASKS["Isaac-Velocity-Flat-Anymal-C-v0"]()
    cfg.observations.policy.joint_pos.history_length = 4
    cfg.observations.policy.joint_vel.modifiers = [modifiers.ModifierCfg(func=modifiers.scale, params={"multiplier": 2.0}),
                                                   modifiers.ModifierCfg(func=modifiers.clip, params={"bounds": (None, 1.5)})]
    spec = compile_env_cfg(cfg, num_envs=4)
    terms = {t["name"]: t for t in spec["observations"]["terms"]}
    assert term_raw_dim(spec, terms["joint_pos"]) == 12 and term_dim(spec, terms["joint_pos"]) == 48
    assert terms["joint_vel"]["modifiers"] == [("scale", 2.0), ("clip", -float("inf"), 1.5)]
    assert sum(term_dim(spec, t) for t in spec["observations"]["terms"]) == 48 + 36
    cfg.observations.policy._group.history_length = 2       # group setting wins over every term's
    spec = compile_env_cfg(cfg, num_envs=4)
    assert all(t["history_length"] == 2 for t in spec["observations"]["terms"])
    assert sum(term_dim(spec, t) for t in spec["observations"]["terms"]) == 96
    cfg.observations.policy._group.flatten_history_dim = False
    with pytest.raises(NotImplementedError, match="flatten_history_dim"):
        compile_env_cfg(cfg, num_envs=4)
    cfg.observations.policy._group.history_length = None

    # a class modifier moves the term to the Python-evaluated (EXTERNAL) path with all its modifiers, in order
    cfg.observations.policy.joint_vel.modifiers = [modifiers.ModifierCfg(func=modifiers.scale, params={"multiplier": 2.0}),
                                                   modifiers.IntegratorCfg(dt=0.1)]
    spec = compile_env_cfg(cfg, num_envs=4)
    jv = {t["name"]: t for t in spec["observations"]["terms"]}["joint_vel"]
    assert jv["op"] == "external" and len(jv["py_modifiers"]) == 2 and "modifiers" not in jv

    def my_modifier(data, gain):
        return data * gain
    cfg.observations.policy.joint_vel.modifiers = [modifiers.ModifierCfg(func=my_modifier, params={"gain": 2.0})]
    spec = compile_env_cfg(cfg, num_envs=4)   # a user modifier function: Python path as well
    jv = {t["name"]: t for t in spec["observations"]["terms"]}["joint_vel"]
    assert jv["op"] == "external" and jv["py_modifiers"][0].func is my_modifier


def test_delayed_pd_cfg_compiles_and_checks_its_range():
    from isaaclab_b200.actuators import DelayedPDActuatorCfg
    from isaaclab_b200.envs.compile import compile_env_cfg
    from isaaclab_b200.tasks import TASKS
    cfg = TASKS["B200-Opcode-Coverage-v0"]()
    spec = compile_env_cfg(cfg, num_envs=4)
    d = [a for a in spec["actuators"] if a["kind"] == "delayed_pd"]
    assert len(d) == 1 and (d[0]["min_delay"], d[0]["max_delay"]) == (1, 3) and len(d[0]["joint_ids"]) == 2
    cfg.scene.robot.actuators["hips_hind"] = DelayedPDActuatorCfg(joint_names_expr=[".*H_HAA"], stiffness=1.0, damping=0.1,
                                                                 min_delay=3, max_delay=1)
    with pytest.raises(ValueError, match="min_delay <= max_delay"):
        compile_env_cfg(cfg, num_envs=4)
