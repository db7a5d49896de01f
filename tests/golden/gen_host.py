"""Golden answers of the reference's name-resolution helpers (LAB/utils/string.py) and observation layout, computed by
the unmodified reference:  python -m tests.golden.gen_host  ->  tests/golden/host_logic.json"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from tests.golden import ref_harness  # noqa: E402

ref_harness.install()
import isaaclab.utils.string as S  # noqa: E402

ANYMAL_J = [f"{l}_{j}" for j in ("HAA", "HFE", "KFE") for l in ("LF", "LH", "RF", "RH")]
ANYMAL_B = ["base"] + [f"{l}_{b}" for b in ("HIP", "THIGH", "SHANK", "FOOT") for l in ("LF", "LH", "RF", "RH")]

NAME_CASES = [
    (".*", ANYMAL_J, False), ([".*HAA", ".*HFE", ".*KFE"], ANYMAL_J, False), ([".*KFE", ".*HAA"], ANYMAL_J, True),
    (".*FOOT", ANYMAL_B, False), (".*THIGH", ANYMAL_B, False), ("base", ANYMAL_B, False),
    (["RH_FOOT", "LF_FOOT"], ANYMAL_B, True), (["RH_FOOT", "LF_FOOT"], ANYMAL_B, False), ("L[F,H]_HIP", ANYMAL_B, False),
]
ERROR_CASES = [(["nope"], ANYMAL_J), ([".*HAA", "LF_.*"], ANYMAL_J)]
VALUE_CASES = [
    ({".*HAA": 0.0, ".*F_HFE": 0.4, ".*H_HFE": -0.4, ".*F_KFE": -0.8, ".*H_KFE": 0.8}, ANYMAL_J, False),
    ({".*": 40.0}, ANYMAL_J, False), ({"LF_KFE": 1.0, ".*HAA": 2.0}, ANYMAL_J, True),
]


def main():
    out = {"names": [], "errors": [], "values": []}
    for keys, strings, order in NAME_CASES:
        idx, names = S.resolve_matching_names(keys, strings, preserve_order=order)
        out["names"].append(dict(keys=keys, strings=strings, preserve_order=order, idx=idx, names=names))
    for keys, strings in ERROR_CASES:
        try:
            S.resolve_matching_names(keys, strings)
            raised = False
        except ValueError:
            raised = True
        out["errors"].append(dict(keys=keys, strings=strings, raises=raised))
    for data, strings, order in VALUE_CASES:
        idx, names, vals = S.resolve_matching_names_values(data, strings, preserve_order=order)
        out["values"].append(dict(data=data, strings=strings, preserve_order=order, idx=idx, names=names, values=vals))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_logic.json")
    json.dump(out, open(path, "w"), indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
