"""Golden vectors of the math helpers the path uses, computed by the UNMODIFIED reference `isaaclab.utils.math`
(LAB/utils/math.py) in the authoring container:  python -m tests.golden.gen_math  ->  tests/golden/math_utils.pt
The wrap_to_pi table repeats the known-answer cases of the reference's own test (LAB/test/utils/test_math.py)."""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from tests.golden import ref_harness  # noqa: E402

ref_harness.install()
import isaaclab.utils.math as M  # noqa: E402


def main():
    g = torch.Generator().manual_seed(11)
    n = 257
    q = torch.randn(n, 4, generator=g)
    q = q / q.norm(dim=-1, keepdim=True)
    q[0] = torch.tensor([1.0, 0.0, 0.0, 0.0])
    q[1] = torch.tensor([0.0, 0.0, 0.0, 1.0])
    q2 = torch.randn(n, 4, generator=g)
    q2 = q2 / q2.norm(dim=-1, keepdim=True)
    v = torch.randn(n, 3, generator=g) * 3.0
    e = (torch.rand(n, 3, generator=g) * 2 - 1) * math.pi
    pi = math.pi
    ang = torch.cat([torch.tensor([0.0, pi, -pi, 2 * pi, -2 * pi, 3 * pi, -3 * pi, 4 * pi, 1e-7, -1e-7, pi + 1e-6, -pi - 1e-6,
                                   10.0, -10.0, 100.5]), (torch.rand(240, generator=g) * 2 - 1) * 25.0])
    out = {
        "in": dict(q=q, q2=q2, v=v, e=e, ang=ang),
        "wrap_to_pi": M.wrap_to_pi(ang),
        "quat_rotate_inverse": M.quat_rotate_inverse(q, v),
        "quat_apply": M.quat_apply(q, v),
        "yaw_quat": M.yaw_quat(q),
        "quat_apply_yaw": M.quat_apply_yaw(q, v),
        "quat_from_euler_xyz": M.quat_from_euler_xyz(e[:, 0], e[:, 1], e[:, 2]),
        "quat_mul": M.quat_mul(q, q2),
        "normalize": M.normalize(v),
        "reference": "isaaclab.utils.math (IsaacLab 2.1.0), torch CPU fp32",
    }
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "math_utils.pt")
    torch.save(out, path)
    print("wrote", path)


if __name__ == "__main__":
    main()
