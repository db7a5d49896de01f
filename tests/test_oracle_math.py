"""oracle/math_utils.py against golden vectors computed by the unmodified reference isaaclab.utils.math
(tests/golden/gen_math.py)."""
import math
import os

import torch

from oracle import math_utils as mu

FX = torch.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "math_utils.pt"), weights_only=False)
I = FX["in"]


def _eq(a, b, name):
    torch.testing.assert_close(a, b, rtol=0, atol=0, msg=lambda m: f"{name}: {m}")


def test_wrap_to_pi_matches_reference_bitwise():
    _eq(mu.wrap_to_pi(I["ang"]), FX["wrap_to_pi"], "wrap_to_pi")


def test_wrap_to_pi_known_answers():
    """Table of LAB/test/utils/test_math.py::test_wrap_to_pi: +pi stays +pi, odd multiples keep their sign."""
    pi = math.pi
    x = torch.tensor([0.0, pi, -pi, 2 * pi, -2 * pi, 3 * pi, -3 * pi, 4 * pi])
    want = torch.tensor([0.0, pi, -pi, 0.0, 0.0, pi, -pi, 0.0])
    torch.testing.assert_close(mu.wrap_to_pi(x), want, rtol=0, atol=1e-6)


def test_quaternion_helpers_match_reference_bitwise():
    q, q2, v, e = I["q"], I["q2"], I["v"], I["e"]
    _eq(mu.quat_rotate_inverse(q, v), FX["quat_rotate_inverse"], "quat_rotate_inverse")
    _eq(mu.quat_apply(q, v), FX["quat_apply"], "quat_apply")
    _eq(mu.yaw_quat(q), FX["yaw_quat"], "yaw_quat")
    _eq(mu.quat_apply_yaw(q, v), FX["quat_apply_yaw"], "quat_apply_yaw")
    _eq(mu.quat_from_euler_xyz(e[:, 0], e[:, 1], e[:, 2]), FX["quat_from_euler_xyz"], "quat_from_euler_xyz")
    _eq(mu.quat_mul(q, q2), FX["quat_mul"], "quat_mul")
    _eq(mu.normalize(v), FX["normalize"], "normalize")


def test_noise_models_match_reference():
    """oracle.mdp_env.apply_noise against the reference noise_model functions (constant / uniform / gaussian x add / scale /
    abs) on the recorded draws -- bitwise."""
    import os
    from oracle.mdp_env import apply_noise
    fx = torch.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "math_utils.pt"), weights_only=False)["noise"]
    assert len(fx["cases"]) == 9
    for c in fx["cases"]:
        got = apply_noise(fx["data"].clone(), c["draw"], (c["kind"], c["lo"], c["hi"], c["op"]))
        assert torch.equal(got, c["out"]), f"{c['kind']}/{c['op']}: max diff {(got - c['out']).abs().max().item()}"
