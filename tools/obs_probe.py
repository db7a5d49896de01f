"""Times b200lab_compute_obs alone on the headline workload (Rough-Anymal-C, 4096 envs, full terrain) -- the per-ray
kernel (B200LAB_OBS_IMPL=0) and the staged one (=1) -- with CUDA events; a short, single-kernel command for ncu.

    python tools/obs_probe.py [--envs 4096] [--iters 200] [--impl 0,1] [--task ...]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=4096)
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--impl", default="0,1")
    ap.add_argument("--task", default="Isaac-Velocity-Rough-Anymal-C-v0")
    ap.add_argument("--no-hits", action="store_true")
    ap.add_argument("--no-noise", action="store_true", help="observation corruption off (no Philox in the kernel)")
    ap.add_argument("--cell", type=float, default=0.0, help="bin-grid cell edge of the terrain mesh (0: the builder's default)")
    args = ap.parse_args()
    from isaaclab_b200 import terrains
    from isaaclab_b200.envs.compile import compile_env_cfg
    from isaaclab_b200.envs.fused_env import FusedEnv
    from isaaclab_b200.sim.synthetic import SyntheticSim, generate_slabs, init_fused_state, initial_env_state
    from isaaclab_b200.tasks import TASKS
    cfg = TASKS[args.task]()
    if args.no_noise:
        cfg.observations.policy._group.enable_corruption = False
    terrain = terrains.TerrainGenerator(cfg.scene.terrain.terrain_generator)
    spec = compile_env_cfg(cfg, num_envs=args.envs, terrain=terrain, seed=0)
    if args.cell > 0:
        spec["raycaster"]["cell_size"] = args.cell
    K = 16
    slabs = generate_slabs(spec, K, seed=1237)
    init = initial_env_state(spec, seed=7)
    out = {}
    for impl in [int(x) for x in args.impl.split(",")]:
        os.environ["B200LAB_OBS_IMPL"] = str(impl)
        fus = FusedEnv(spec, device="cuda", seed=0, store_ray_hits=not args.no_hits)
        init_fused_state(fus, init)
        sim = SyntheticSim(spec, K=K, device="cuda", slabs={k: v.clone() for k, v in slabs.items()})
        action = torch.randn(args.envs, spec["A"], device="cuda")
        for k in range(4):   # real steps: drift, resets, commands in a realistic state
            fus.step(action, sim.views(k + 1), sim.views(k)[-1])
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.iters)]
        for i in range(args.iters + 5):
            fus.rc_is_outdated.fill_(1)
            view = sim.views(i)[-1]
            if i >= 5:
                ev[i - 5][0].record()
            fus.compute_obs(view)
            if i >= 5:
                ev[i - 5][1].record()
        torch.cuda.synchronize()
        ts = sorted(a.elapsed_time(b) * 1e3 for a, b in ev)
        out[impl] = (ts[len(ts) // 2], ts[0])
        print(f"impl {impl}: compute_obs (derive + obs kernels) median {ts[len(ts) // 2]:.1f} us, best {ts[0]:.1f} us", flush=True)
        del fus, sim
    return out


if __name__ == "__main__":
    main()
