"""Ray patterns against the reference's own functions (tests/golden/gen_patterns.py ran them from /root/reference)."""
import os

import pytest
import torch

from isaaclab_b200.sensors import patterns as P

_CFG = {"grid": P.GridPatternCfg, "lidar": P.LidarPatternCfg, "bpearl": P.BpearlPatternCfg, "pinhole": P.PinholeCameraPatternCfg}


def test_patterns_match_reference(golden_dir):
    fx = torch.load(os.path.join(golden_dir, "patterns.pt"), weights_only=False)
    for name, (kind, kw) in fx["cases"].items():
        cfg = _CFG[kind](**kw)
        s, d = cfg.func(cfg, fx["intrinsics"], "cpu") if kind == "pinhole" else cfg.func(cfg, "cpu")
        rs, rd = fx[name]
        assert s.shape == rs.shape and d.shape == rd.shape, name
        torch.testing.assert_close(s, rs, rtol=0, atol=1e-7, msg=lambda m: f"{name} starts: {m}")   # same ray ORDER
        torch.testing.assert_close(d, rd, rtol=1e-6, atol=1e-7, msg=lambda m: f"{name} directions: {m}")


def test_pinhole_cfg_from_intrinsic_matrix():
    c = P.PinholeCameraPatternCfg.from_intrinsic_matrix([8.0, 0.0, 3.5, 0.0, 9.0, 2.5, 0.0, 0.0, 1.0], width=7, height=5)
    assert c.horizontal_aperture == pytest.approx(7 * 24.0 / 8.0) and c.vertical_aperture == pytest.approx(5 * 24.0 / 9.0)
    assert c.horizontal_aperture_offset == pytest.approx(0.0) and c.vertical_aperture_offset == pytest.approx(0.0)
