"""Would running the step on env halves in parallel streams pay?  P independent FusedEnvs of 4096/P envs each (same task,
full terrain), each replaying its own captured step graphs on its own stream, free-running; total env-steps/s against one
4096-env instance.  usage: python tools/split_stream_probe.py [steps=400]"""
import gc
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from isaaclab_b200.envs.fused_env import FusedEnv  # noqa: E402
from isaaclab_b200.sim.synthetic import SyntheticSim, init_fused_state  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 400
dev = torch.device("cuda", 0)


def build(envs, seed):
    cfg, spec, slabs, init = bench.build_workload("cfg3", seed=seed, device=dev, envs=envs)
    K = slabs["root_state_w"].shape[0]
    sim = SyntheticSim(spec, K=K, device=dev, slabs=slabs)
    fus = FusedEnv(spec, device=dev, seed=seed, store_ray_hits=False)
    init_fused_state(fus, init)
    action = torch.randn(envs, spec["A"], generator=torch.Generator().manual_seed(seed)).to(dev)
    for k in range(3):
        fus.step(action, sim.views(k + 1), sim.views(k)[-1])
    fus.use_device_step_counter()
    graphs = []
    gc.disable()
    try:
        for j in range(K):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, capture_error_mode="thread_local"):
                fus.step(action, sim.views(j + 1), sim.views(j)[-1])
            graphs.append(gr)
    finally:
        gc.enable()
    return fus, sim, action, graphs


for P in (1, 2, 4):
    parts = [build(4096 // P, 10 + i) for i in range(P)]
    streams = [torch.cuda.Stream() for _ in range(P)]
    K = len(parts[0][3])
    for k in range(10):
        for p, st in zip(parts, streams):
            with torch.cuda.stream(st):
                p[3][k % K].replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for st in streams:
        st.wait_event(e0)
    for k in range(steps):
        for p, st in zip(parts, streams):
            with torch.cuda.stream(st):
                p[3][(10 + k) % K].replay()
    for st in streams:
        torch.cuda.current_stream().wait_stream(st)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"{P} x {4096 // P} envs on {P} stream(s): {ms / steps * 1e3:7.1f} us per 4096 env-steps, {4096 * steps / ms / 1e3:6.2f} M env-steps/s", flush=True)
    del parts, streams
    torch.cuda.empty_cache()
