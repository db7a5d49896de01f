"""tests/golden/patterns.pt: the reference's ray patterns (LAB/sensors/ray_caster/patterns/patterns.py) evaluated here from
/root/reference through the stub loader.  Run in the authoring container only: python tests/golden/gen_patterns.py"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness  # noqa: E402

ref_harness.install()
from isaaclab.sensors.ray_caster.patterns import patterns, patterns_cfg  # noqa: E402

CASES = {
    "grid_xy": ("grid", dict(resolution=0.1, size=(1.6, 1.0), direction=(0.0, 0.0, -1.0), ordering="xy")),
    "grid_yx": ("grid", dict(resolution=0.25, size=(1.0, 2.0), direction=(0.0, 0.3, -1.0), ordering="yx")),
    "lidar_360": ("lidar", dict(channels=8, vertical_fov_range=(-15.0, 15.0), horizontal_fov_range=(-180.0, 180.0), horizontal_res=5.0)),
    "lidar_front": ("lidar", dict(channels=3, vertical_fov_range=(-5.0, 20.0), horizontal_fov_range=(-45.0, 60.0), horizontal_res=7.0)),
    "bpearl": ("bpearl", dict()),
    "bpearl_coarse": ("bpearl", dict(horizontal_fov=180.0, horizontal_res=22.5, vertical_ray_angles=[80.0, 45.0, 10.0])),
    "pinhole": ("pinhole", dict(width=7, height=5)),
}
INTRINSICS = torch.tensor([[[8.0, 0.0, 3.5], [0.0, 8.0, 2.5], [0.0, 0.0, 1.0]], [[6.0, 0.0, 3.0], [0.0, 9.0, 2.0], [0.0, 0.0, 1.0]]])

if __name__ == "__main__":
    out = {"cases": CASES, "intrinsics": INTRINSICS}
    for name, (kind, kw) in CASES.items():
        if kind == "grid":
            s, d = patterns.grid_pattern(patterns_cfg.GridPatternCfg(**kw), "cpu")
        elif kind == "lidar":
            s, d = patterns.lidar_pattern(patterns_cfg.LidarPatternCfg(**kw), "cpu")
        elif kind == "bpearl":
            s, d = patterns.bpearl_pattern(patterns_cfg.BpearlPatternCfg(**kw), "cpu")
        else:
            s, d = patterns.pinhole_camera_pattern(patterns_cfg.PinholeCameraPatternCfg(**kw), INTRINSICS, "cpu")
        out[name] = (s.clone(), d.clone())
        print(name, tuple(s.shape))
    torch.save(out, os.path.join(HERE, "patterns.pt"))
