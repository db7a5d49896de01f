"""Every task cfg of isaaclab_b200.tasks against the numbers of the reference's own cfg objects
(tests/golden/task_cfgs.json, written by tests/golden/gen_task_cfgs.py from /root/reference): reward / termination /
observation / event terms with weights and params, action scale, command ranges, terrain overrides, actuator models."""
import importlib.util
import json
import os

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = json.load(open(os.path.join(HERE, "golden", "task_cfgs.json")))


def _digest_fn():
    """The same digest function that produced the fixture, without running the generator's reference import."""
    src = open(os.path.join(HERE, "golden", "gen_task_cfgs.py")).read()
    body = src[src.index("def _plain"):src.index('if __name__ == "__main__":')]
    ns = {}
    exec(body, ns)
    return ns["digest"]


def _norm(x):
    """Differences that are representation only: tuples / lists, ints / floats, regex lists of one element."""
    if isinstance(x, dict):
        return {k: _norm(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return [_norm(v) for v in x]
    if isinstance(x, bool) or x is None or isinstance(x, str):
        return x
    if isinstance(x, (int, float)):
        return round(float(x), 9)
    return x


# fields that exist only on one side of the boundary (USD / PhysX / visualisation), by section
_SKIP_EVENTS = ("physics_material", "add_base_mass", "base_com", "base_external_force_torque")


@pytest.mark.parametrize("task", sorted(GOLDEN))
def test_task_cfg_matches_reference(task):
    from isaaclab_b200.tasks import TASKS
    ref = _norm(GOLDEN[task])
    got = _norm(_digest_fn()(TASKS[task]()))
    for key in ("decimation", "sim_dt", "episode_length_s", "num_envs", "env_spacing", "enable_corruption", "commands",
                "height_scanner"):
        assert got[key] == ref[key], f"{task}: {key}: {got[key]} != {ref[key]}"
    for section in ("rewards", "terminations", "observations", "curriculum"):
        assert list(got[section]) == list(ref[section]), f"{task}: {section} term names / order"
        for name, r in ref[section].items():
            g = got[section][name]
            for f in ("func", "weight", "time_out", "scale", "clip", "noise", "history_length"):
                if f in r:
                    assert g.get(f) == r[f], f"{task}: {section}.{name}.{f}: {g.get(f)} != {r[f]}"
            for pk, pv in r["params"].items():
                assert g["params"].get(pk) == pv, f"{task}: {section}.{name}.params[{pk}]: {g['params'].get(pk)} != {pv}"
    ev_ref = {k: v for k, v in ref["events"].items() if k not in _SKIP_EVENTS}
    ev_got = {k: v for k, v in got["events"].items() if k not in _SKIP_EVENTS}
    assert list(ev_got) == list(ev_ref), f"{task}: event terms"
    for name, r in ev_ref.items():
        assert ev_got[name]["func"] == r["func"] and ev_got[name]["mode"] == r["mode"], f"{task}: events.{name}"
        assert ev_got[name].get("interval_range_s") == r.get("interval_range_s"), f"{task}: events.{name}.interval_range_s"
        for pk, pv in r["params"].items():
            if pk != "asset_cfg":
                assert ev_got[name]["params"].get(pk) == pv, f"{task}: events.{name}.params[{pk}]"
    for name, r in ref["actions"].items():
        assert got["actions"][name]["scale"] == r["scale"] and got["actions"][name]["joint_names"] == r["joint_names"], f"{task}: actions.{name}"
    if ref["terrain"] is not None:
        for k, v in ref["terrain"].items():
            assert got["terrain"][k] == v, f"{task}: terrain.{k}: {got['terrain'][k]} != {v}"
    assert got["robot"]["init_pos"] == ref["robot"]["init_pos"] and got["robot"]["joint_pos"] == ref["robot"]["joint_pos"], f"{task}: robot init"
    assert got["robot"]["soft_joint_pos_limit_factor"] == ref["robot"]["soft_joint_pos_limit_factor"]
    assert list(got["robot"]["actuators"]) == list(ref["robot"]["actuators"]), f"{task}: actuator groups"
    for name, r in ref["robot"]["actuators"].items():
        g = got["robot"]["actuators"][name]
        for f, v in r.items():
            assert g.get(f) == v, f"{task}: actuators.{name}.{f}: {g.get(f)} != {v}"
