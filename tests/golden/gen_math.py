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


def noise_vectors(g):
    """constant / uniform / gaussian noise x add / scale / abs by the reference's noise_model functions
    (LAB/utils/noise/noise_model.py:20-98) with the rand_like / randn_like draws recorded."""
    from isaaclab.utils.noise import noise_cfg, noise_model
    data = torch.randn(64, 5, generator=g) * 2.0
    draws = {}
    orig_rand, orig_randn = torch.rand_like, torch.randn_like

    def rand_like(x, **k):
        draws["u"] = torch.rand(x.shape, generator=g)
        return draws["u"]

    def randn_like(x, **k):
        draws["u"] = torch.randn(x.shape, generator=g)
        return draws["u"]

    cases = []
    torch.rand_like, torch.randn_like = rand_like, randn_like
    try:
        for op in ("add", "scale", "abs"):
            for kind, cfg, fn, lo, hi in (
                    ("constant", noise_cfg.ConstantNoiseCfg(bias=0.37, operation=op), noise_model.constant_noise, 0.37, 0.0),
                    ("uniform", noise_cfg.UniformNoiseCfg(n_min=-0.1, n_max=0.25, operation=op), noise_model.uniform_noise, -0.1, 0.25),
                    ("gaussian", noise_cfg.GaussianNoiseCfg(mean=0.05, std=0.3, operation=op), noise_model.gaussian_noise, 0.05, 0.3)):
                draws.clear()
                y = fn(data.clone(), cfg)
                cases.append(dict(kind=kind, op=op, lo=lo, hi=hi, draw=draws.get("u", torch.zeros_like(data)).clone(), out=y.clone()))
    finally:
        torch.rand_like, torch.randn_like = orig_rand, orig_randn
    return dict(data=data, cases=cases)


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
    out["noise"] = noise_vectors(g)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "math_utils.pt")
    torch.save(out, path)
    print("wrote", path)


if __name__ == "__main__":
    main()
