"""Runs the fused env step of one BASELINE config eagerly (short, for ncu):  python tools/step_probe.py --cfg cfg4 --iters 10"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", default="cfg4")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--live", action="store_true")
    a = ap.parse_args()
    from isaaclab_b200.envs.fused_env import FusedEnv
    from isaaclab_b200.sim.synthetic import SyntheticSim, init_fused_state
    dev = torch.device("cuda", 0)
    cfg, spec, slabs, init = bench.build_workload(a.cfg, seed=0, device=dev)
    sim = SyntheticSim(spec, K=slabs["root_state_w"].shape[0], device=dev, slabs=slabs)
    fus = FusedEnv(spec, device=dev, seed=0, store_ray_hits=False)
    init_fused_state(fus, init)
    action = torch.randn(spec["N"], spec["A"], device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for k in range(a.iters):
        if k == a.iters // 2:
            e0.record()
        fus.step(action, sim.views(k + 1), sim.views(k)[-1], fused=not a.live)
    e1.record()
    torch.cuda.synchronize()
    print(f"{a.cfg}: {e0.elapsed_time(e1) / (a.iters - a.iters // 2) * 1e3:.1f} us / step")


if __name__ == "__main__":
    main()
