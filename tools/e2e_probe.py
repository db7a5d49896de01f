"""Where does the end-to-end step time go?  Times ManagerBasedRLEnv.step variants on the bench workload:
host/device simulator state, eager/graph, with/without result read-back; prints wall us/step and the host time spent
inside step() (enqueue cost).  usage: python tools/e2e_probe.py [steps]"""
import gc
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from isaaclab_b200.envs.manager_based_rl_env import DeviceReplaySim, HostReplaySim, ManagerBasedRLEnv  # noqa: E402
from isaaclab_b200.sim.synthetic import SyntheticSim  # noqa: E402
from tests.mdp_common import init_fused_state  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
dev = torch.device("cuda", 0)
cfg, spec, slabs, init = bench.build_workload("cfg3", seed=0, device=dev)
N = spec["N"]
g = torch.Generator().manual_seed(1)
act_host = torch.randn(N, spec["A"], generator=g).pin_memory()
act_dev = act_host.to(dev)


def run(name, sim_kind, graph, results, prefetch=True, chunks=4):
    if sim_kind == "host":
        sim = HostReplaySim(spec, slabs, device=dev, prefetch=prefetch)
    else:
        sim = DeviceReplaySim(SyntheticSim(spec, K=slabs["root_state_w"].shape[0], device=dev, slabs=slabs))
    env = ManagerBasedRLEnv(cfg, sim=sim, device=dev, spec=spec, seed=0, use_cuda_graph=graph)
    env.result_chunks = chunks
    init_fused_state(env.fused, init)
    res = torch.empty(env.fused.result.numel(), dtype=torch.uint8).pin_memory() if results else None
    act = act_host if sim_kind == "host" or results else act_dev
    host_t = 0.0

    def one():
        nonlocal host_t
        t0 = time.perf_counter()
        env.step(act, results_out=res)
        host_t += time.perf_counter() - t0
        if res is not None:
            env.wait_results()
        else:
            torch.cuda.current_stream().synchronize()

    for _ in range(40):
        one()
    torch.cuda.synchronize()
    host_t = 0.0
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / steps * 1e6
    print(f"{name:58s} {wall:7.1f} us/step   host time in step() {host_t / steps * 1e6:6.1f} us", flush=True)
    env.close()
    del env, sim
    gc.collect()


run("device sim, eager, no read-back, sync/step", "dev", False, False)
run("device sim, graph, no read-back, sync/step", "dev", True, False)
run("device sim, eager, read-back (1 copy)", "dev", False, True)
run("device sim, graph, read-back pipelined x4", "dev", True, True)
run("device sim, graph, read-back pipelined x1", "dev", True, True, chunks=1)
run("host sim (prefetch), eager, read-back (1 copy)", "host", False, True)
run("host sim (prefetch), graph, read-back pipelined x4", "host", True, True)
run("host sim (prefetch), graph, read-back pipelined x2", "host", True, True, chunks=2)
run("host sim (prefetch), graph, read-back pipelined x1", "host", True, True, chunks=1)
run("host sim (no prefetch), graph, read-back pipelined x4", "host", True, True, prefetch=False)
