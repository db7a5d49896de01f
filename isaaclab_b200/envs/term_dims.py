"""Width of each observation term (ObservationManager._prepare_terms infers it by calling the term once,
LAB/managers/observation_manager.py:420-450; here it follows from the term identity)."""


def term_dim(spec, t):
    """Width in the concatenated group: the term's own width times its flattened history length
    (observation_manager.py:422-433)."""
    return term_raw_dim(spec, t) * max(1, int(t.get("history_length", 0) or 0))


def term_raw_dim(spec, t):
    op = t["op"]
    if op in ("base_lin_vel", "base_ang_vel", "projected_gravity", "generated_commands"):
        return 3
    if op in ("joint_pos_rel", "joint_vel_rel", "joint_pos", "joint_vel"):
        return len(t["joint_ids"]) if t.get("joint_ids") is not None else spec["J"]
    if op == "last_action":
        return spec["A"]
    if op == "height_scan":
        return spec["R"]
    if op == "base_pos_z":
        return 1
    if op == "external":
        if t.get("dim") is None:
            raise ValueError(f"observation term '{t.get('name')}': the width of a user term is set when the env is built")
        return int(t["dim"])
    raise ValueError(f"unknown observation term '{op}'")
