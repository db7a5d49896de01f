"""How much of the observation / ray-cast kernel can hide under the decimation + post_step kernels when launched on a
second stream?  (Same work as a real step; ordering semantics ignored -- this only measures achievable overlap.)"""
import sys
sys.path.insert(0, ".")
import torch
import bench
from isaaclab_b200.envs.fused_env import FusedEnv
from isaaclab_b200.sim.synthetic import SyntheticSim
from tests.mdp_common import init_fused_state

dev = torch.device("cuda", 0)
cfg, spec, slabs, init = bench.build_workload(dev, 0)
sim = SyntheticSim(spec, K=bench.K_SLABS, device=dev, slabs=slabs)
fus = FusedEnv(spec, device=dev, seed=0)
init_fused_state(fus, init)
N = spec["N"]
action = torch.randn(N, spec["A"], device=dev)
side = torch.cuda.Stream()
main = torch.cuda.current_stream()


def seq(k):
    fus.step(action, sim.views(k + 1), sim.views(k)[-1])


def overlap(k, with_decimation):
    views, prev = sim.views(k + 1), sim.views(k)[-1]
    ev0 = torch.cuda.Event(); ev1 = torch.cuda.Event(); ev2 = torch.cuda.Event()
    if with_decimation:
        ev0.record(main); side.wait_event(ev0)
        with torch.cuda.stream(side):
            fus.compute_obs(views[-1]); ev1.record(side)
        fus.decimation_fused(action, views, prev)
        fus.post_step(views[-1])
    else:
        fus.decimation_fused(action, views, prev)
        ev0.record(main); side.wait_event(ev0)
        with torch.cuda.stream(side):
            fus.compute_obs(views[-1]); ev1.record(side)
        fus.post_step(views[-1])
    fus.post_reset_write(views[-1])
    main.wait_event(ev1)
    fus.step_counter += 1


def timeit(fn, steps=300, warm=20):
    for k in range(warm):
        fn(k)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(steps):
        fn(warm + k)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps * 1e3


print("sequential step           %.1f us" % timeit(seq))
print("obs || post_step          %.1f us" % timeit(lambda k: overlap(k, False)))
print("obs || decimation+post    %.1f us" % timeit(lambda k: overlap(k, True)))
