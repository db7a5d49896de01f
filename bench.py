"""bench.py -- headline benchmark of the B200-native non-physics step (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one ManagerBasedRLEnv.step worth of non-physics work over the env shard of one GPU:
process_action + S x (actuator model, contact/joint-acc update) + terminations/rewards/resets/commands/events +
observations incl. the 187-ray height scan -- on synthetic ArticulationData-shaped state (PhysX is out of scope).
Workload at N=1: BASELINE.json configs[2] "Isaac-Velocity-Rough-Anymal-C-v0, 4096 envs/GPU" (the config the metric is
quoted on).  Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

TASK = "Isaac-Velocity-Rough-Anymal-C-v0"
ENVS_PER_GPU = 4096
K_SLABS = 32
METRIC = "non-physics env-steps/s (Anymal-C rough, 4096 envs/GPU)"
UNIT = "env-steps/s"


def algorithmic_bytes(spec):
    """Compulsory HBM bytes per env per launch of each kernel (split per kernel)."""
    J, A, B, S = spec["J"], spec["A"], spec["B"], spec["decimation"]
    F = 4
    Tr = len(spec["rewards"])
    state = 0
    for a in spec["actuators"]:
        Jg = len(a["joint_ids"])
        if a["kind"] == "lstm":
            state += 16 * a["lstm"]["num_layers"] * Jg * a["lstm"]["hidden"]
        elif a["kind"] == "mlp":
            state += 16 * Jg + 8 * (max(a["mlp"]["input_idx"]) + 1) * Jg
    obs = sum(_term_dim(spec, t) for t in spec["observations"]["terms"])
    per = {
        "process_action": 24 * A,
        "substep_pre": 20 * J + state,
        "substep_post": (12 * B + 12 * B + 32 * B if spec.get("contact") else 0) + 8 * J,
        "post_step": 52 + 12 * J + 4 * J + 8 * A + 36 * B + 8 * F + 8 + 30 + 4 + 2 + 12 * Tr + 16 + 8 + 4 * J,
        "compute_obs": (28 if spec.get("raycaster") else 0) + 4 * obs,
    }
    per["decimation_fused"] = S * (per["substep_pre"] + per["substep_post"])
    # actuator pass of the write_data_to_sim() that follows _reset_idx (runs whenever any env reset: at 4096 envs, always)
    per["post_reset_write"] = per["substep_pre"]
    per_step = per["decimation_fused"] + per["post_step"] + per["post_reset_write"] + per["compute_obs"]
    return per, per_step


def _term_dim(spec, t):
    from isaaclab_b200.envs.term_dims import term_dim
    return term_dim(spec, t)


class ClockSampler(threading.Thread):
    """SM clock and clock-event (throttle) reasons sampled DURING the timed region: NVML polled every 2 ms (the timed
    region of a default run is well under a second -- nvidia-smi -lms cannot sample that), nvidia-smi as a fallback
    (B200_PROFILING.md clocks line)."""

    _REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
                ("sw_power_cap", 0x4))

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False
        self.max_mhz, self.mask, self.proc, self.source = None, 0, None, "nvml"
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                self.index = int(vis.split(",")[index])
            except (ValueError, IndexError):
                pass

    def _run_nvml(self):
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(self.index)
        self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag:
            self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
            self.mask |= int(reasons(h))
            time.sleep(0.002)

    def _run_smi(self):
        self.source = "nvidia-smi"
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index), "-lms", "20"], stdout=subprocess.PIPE, text=True)
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            if f and f[0].isdigit():
                self.samples.append(int(f[0]))
                if len(f) > 1 and f[1].isdigit():
                    self.max_mhz = int(f[1])
                for (name, bit), v in zip(self._REASONS, f[2:6]):
                    if v.lower().startswith("active"):
                        self.mask |= bit
            if self.stop_flag:
                break

    def run(self):
        try:
            self._run_nvml()
        except Exception:
            try:
                self._run_smi()
            except Exception:
                pass

    def finish(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2.0)
        sm = sorted(self.samples)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": [name for name, bit in self._REASONS if self.mask & bit], "samples": len(sm),
                "source": self.source}


_UNIT_BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
_NCU_KERNEL = {"compute_obs": "compute_obs_kernel", "decimation_fused": "decimation_fused_kernel",
               "post_step": "post_step_kernel", "post_reset_write": "substep_pre_kernel"}


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed `ncu --set full` summary
    (profiles/*_ncu_step_kernels.json, tools/ncu_summary.py); None when no capture is committed."""
    import glob
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_step_kernels.json")), reverse=True):
        for recs in json.load(open(path)).values():
            for r in recs:
                if _NCU_KERNEL.get(kernel, kernel) in r["kernel"] and "dram_read" in r:
                    return int(r["dram_read"] * _UNIT_BYTES[r.get("dram_read_unit", "byte")]
                               + r.get("dram_write", 0.0) * _UNIT_BYTES[r.get("dram_write_unit", "byte")])
    return None


def workload_name(spec):
    return (f"{TASK} non-physics step, {spec['N']} envs/GPU, 187-ray height scan, ActuatorNetLSTM, "
            f"decimation {spec['decimation']}")


def build_workload(device, seed):
    from isaaclab_b200 import terrains
    from isaaclab_b200.envs.compile import compile_env_cfg
    from isaaclab_b200.sim.synthetic import generate_slabs, initial_env_state
    from isaaclab_b200.tasks import TASKS
    cfg = TASKS[TASK]()
    terrain = terrains.TerrainGenerator(cfg.scene.terrain.terrain_generator)
    spec = compile_env_cfg(cfg, num_envs=ENVS_PER_GPU, terrain=terrain, seed=seed)
    slabs = generate_slabs(spec, K_SLABS, seed=1234 + 3 + seed)
    init = initial_env_state(spec, seed=7 + seed)
    return cfg, spec, slabs, init


def time_kernels(fus, sim, action, steps):
    """Per-kernel CUDA-event durations over `steps` env steps on the launching stream (ms per launch, averaged).
    Replay path: decimation_fused (process_action + S x (pre, post)), post_step, post_reset_write (conditional actuator
    pass), compute_obs (derive + obs kernels)."""
    names = ["decimation_fused", "post_step", "post_reset_write", "compute_obs"]
    ev = {n: [] for n in names}

    def rec(name, fn, *a):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(*a)
        e1.record()
        ev[name].append((e0, e1))

    k0 = 1000
    for k in range(steps):
        prev = sim.views(k0 + k)[-1]
        views = sim.views(k0 + k + 1)
        rec("decimation_fused", fus.decimation_fused, action, views, prev)
        rec("post_step", fus.post_step, views[-1])
        rec("post_reset_write", fus.post_reset_write, views[-1])
        rec("compute_obs", fus.compute_obs, views[-1])
        fus.step_counter += 1
    torch.cuda.synchronize()
    return {n: sum(a.elapsed_time(b) for a, b in ev[n]) / max(1, len(ev[n])) for n in names}


def run_ours(args):
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from isaaclab_b200 import _lib
    from isaaclab_b200.envs.fused_env import FusedEnv
    from isaaclab_b200.envs.manager_based_rl_env import HostReplaySim, ManagerBasedRLEnv
    from isaaclab_b200.sim.synthetic import SyntheticSim
    from isaaclab_b200.sim.synthetic import init_fused_state

    cfg, spec, slabs, init = build_workload(dev, seed=rank)
    N = spec["N"]
    L = _lib.lib()
    # ---- device-resident arm: inputs in HBM, K slabs (> L2) cycled
    sim = SyntheticSim(spec, K=K_SLABS, device=dev, slabs=slabs)
    fus = FusedEnv(spec, device=dev, seed=rank)
    init_fused_state(fus, init)
    g = torch.Generator().manual_seed(11 + rank)
    action = torch.randn(N, spec["A"], generator=g).to(dev)

    def one_step(k):
        fus.step(action, sim.views(k + 1), sim.views(k)[-1])

    for k in range(args.warmup):
        one_step(k)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    torch.cuda.synchronize()
    launches0 = L.b200lab_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(args.steps):
        one_step(args.warmup + k)
    e1.record()
    torch.cuda.synchronize()
    launches = L.b200lab_launch_count() - launches0
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.barrier()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    clocks = sampler.finish()
    value = world * N * args.steps / (ms * 1e-3)
    # ---- per-kernel timing + roofline of the dominant kernel
    per_kernel_ms = time_kernels(fus, sim, action, max(8, min(args.steps, 64)))
    bytes_per, bytes_step = algorithmic_bytes(spec)
    S = spec["decimation"]
    dom = max(per_kernel_ms, key=per_kernel_ms.get)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = bytes_per[dom] * N / (per_kernel_ms[dom] * 1e-3) / 1e9
    roof = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
            "frac": round(achieved / peak, 4), "traffic": ncu_traffic(dom), "peak_source": "measured" if peaks else "fallback",
            "algorithmic_bytes_per_launch": bytes_per[dom] * N,
            "per_kernel_us": {n: round(v * 1e3, 2) for n, v in per_kernel_ms.items()},
            "step_algorithmic_bytes": bytes_step * N,
            "step_frac": round(bytes_step * N / (ms / args.steps * 1e-3) / 1e9 / peak, 4)}
    # ---- end-to-end arm: public API with HOST buffers (pinned), H2D of each step's simulator state + action and
    # D2H of obs / reward / dones inside the timed region
    host_sim = HostReplaySim(spec, slabs, device=dev)
    env = ManagerBasedRLEnv(cfg, sim=host_sim, device=dev, spec=spec, seed=rank, use_cuda_graph=True)
    init_fused_state(env.fused, init)
    act_host = torch.randn(N, spec["A"], generator=g).pin_memory()
    env.result_chunks = 4
    res_host = torch.empty(env.fused.result.numel(), dtype=torch.uint8).pin_memory()

    def e2e_step():
        # obs, reward, terminated, time-outs, reset mask stream into the pinned slab behind the kernels that produce them
        env.step(act_host, results_out=res_host)
        env.wait_results()  # the caller consumes the result before the next step

    for _ in range(args.warmup):
        e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    torch.cuda.synchronize()
    ms_e2e = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms_e2e, op=dist.ReduceOp.MAX)
    ms_e2e = float(ms_e2e.item())
    h2d = host_sim.h2d_bytes_per_step + act_host.numel() * 4
    d2h = res_host.numel()
    e2e = {"value": round(world * N * args.steps / (ms_e2e * 1e-3), 1), "unit": UNIT, "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "ms_per_step": round(ms_e2e / args.steps, 4)}
    out = {"metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 5), "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": workload_name(spec), "envs_per_gpu": N,
                      "e2e_mode": "ManagerBasedRLEnv.step(action, results_out) with pinned host buffers, the step captured "
                                  "as a CUDA graph per staging set (H2D action, kernels, D2H of reward/masks after "
                                  "post_step and of the obs rows range by range behind the obs kernels); results "
                                  "synchronised every step; next step's simulator state prefetched H2D on a copy "
                                  "stream while the current step computes", "l2_policy": f"{K_SLABS} state slabs "
                                  f"({sim.bytes() >> 20} MiB > 126 MiB L2) cycled", "parallelism": f"env-shard x{world}"},
           "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roof}
    del env, host_sim, fus, sim
    torch.cuda.empty_cache()
    mg = {"global_rank": rank, "local_rank": local, "world_size": world} if world > 1 else None
    out["ppo_update"] = time_ppo_update(dev, precision=2, multi_gpu_cfg=mg)   # N>1: one NCCL all-reduce per minibatch
    out["ppo_update"]["fp32_ffma_s_per_iter"] = time_ppo_update(dev, precision=0, multi_gpu_cfg=mg)["s_per_iter"]
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(spec, slabs, init, budget_s=15.0)
            out["cpu_baseline"]["ppo_update"] = cpu_ppo_baseline()
    if world > 1:
        dist.destroy_process_group()
    return json.dumps(out) if rank == 0 else None


PPO_CFG = dict(N=ENVS_PER_GPU, T=24, obs_dim=235, act_dim=12, hidden=[512, 256, 128], epochs=5, mini_batches=4, clip_param=0.2,
               gamma=0.99, lam=0.95, value_loss_coef=1.0, entropy_coef=0.005, learning_rate=1e-3, max_grad_norm=1.0,
               desired_kl=0.01)   # config 3: TASKS/.../anymal_c/agents/rsl_rl_ppo_cfg.py (AnymalCRoughPPORunnerCfg)


def synthetic_rollout(c, seed, forward=None):
    """SURVEY 8d rollout: obs N(0,1), rewards N(0,0.05), dones Bernoulli(0.01), sigma 1.  With `forward`
    (obs[M,D] -> (mu[M,A], value[M,1]), the arm's own random-init policy) the rollout is what `PPO.act` would have
    stored: mu/values from the policy, actions = mu + sigma*eps, log-prob of those actions -- so the first minibatch
    starts at ratio 1 as in training.  Without it mu/values are N(0,1)."""
    g = torch.Generator().manual_seed(seed)
    T, N, D, A = c["T"], c["N"], c["obs_dim"], c["act_dim"]
    r = dict(observations=torch.randn(T, N, D, generator=g), sigma=torch.ones(T, N, A),
             rewards=0.05 * torch.randn(T, N, 1, generator=g), dones=(torch.rand(T, N, 1, generator=g) < 0.01).to(torch.uint8))
    eps = torch.randn(T, N, A, generator=g)
    r["last_values"] = torch.randn(N, 1, generator=g)
    if forward is None:
        r["mu"], r["values"] = torch.randn(T, N, A, generator=g), torch.randn(T, N, 1, generator=g)
    else:
        mu, val = zip(*(forward(r["observations"][t]) for t in range(T)))
        r["mu"], r["values"] = torch.stack([m.cpu() for m in mu]), torch.stack([v.cpu() for v in val])
    r["actions"] = r["mu"] + r["sigma"] * eps
    r["actions_log_prob"] = (-0.5 * ((r["actions"] - r["mu"]) / r["sigma"]) ** 2 - torch.log(r["sigma"])
                             - 0.5 * math.log(2 * math.pi)).sum(-1, keepdim=True)
    return r


def ppo_flops_per_iter(c):
    dims = [c["obs_dim"], *c["hidden"]]
    fwd = 2 * (sum(a * b for a, b in zip(dims[:-1], dims[1:])) * 2 + c["hidden"][-1] * (c["act_dim"] + 1))
    return 3.0 * fwd * c["T"] * c["N"] * c["epochs"]   # fwd + dX + dW (upper bound)


def time_ppo_update(dev, reps=3, precision=2, multi_gpu_cfg=None):
    """Second half of the metric: PPO update s/iter = GAE + 5 epochs x 4 minibatches (gather, actor/critic fwd, fused
    loss fwd+bwd, bwd GEMMs, KL-adaptive LR, clip + Adam) at config-3 shapes, CUDA events on the launching stream."""
    from isaaclab_b200.rsl_rl import PPO, ActorCritic, TensorDict
    c = PPO_CFG
    torch.manual_seed(1234 + 3)
    obs0 = TensorDict({"policy": torch.zeros(c["N"], c["obs_dim"], device=dev)}, batch_size=[c["N"]])
    pol = ActorCritic(obs0, {"policy": ["policy"], "critic": ["policy"]}, c["act_dim"], actor_hidden_dims=c["hidden"],
                      critic_hidden_dims=c["hidden"], device=dev, precision=precision)
    alg = PPO(pol, num_learning_epochs=c["epochs"], num_mini_batches=c["mini_batches"], clip_param=c["clip_param"],
              gamma=c["gamma"], lam=c["lam"], value_loss_coef=c["value_loss_coef"], entropy_coef=c["entropy_coef"],
              learning_rate=c["learning_rate"], max_grad_norm=c["max_grad_norm"], desired_kl=c["desired_kl"], device=dev,
              multi_gpu_cfg=multi_gpu_cfg)
    if multi_gpu_cfg is not None:
        alg.broadcast_parameters()
    alg.init_storage("rl", c["N"], c["T"], obs0, [c["act_dim"]])
    def forward(o):
        od = TensorDict({"policy": o.to(dev)}, batch_size=[o.shape[0]])
        return pol.act_inference(od).clone(), pol.evaluate(od).clone()

    roll = {k: v.to(dev) for k, v in synthetic_rollout(c, 1234 + 3, forward).items()}
    st = alg.storage
    L = __import__("isaaclab_b200._lib", fromlist=["lib"]).lib()

    def load():
        st.observations["policy"].copy_(roll["observations"])
        for k in ("actions", "rewards", "values", "actions_log_prob", "mu", "sigma", "dones"):
            getattr(st, k).copy_(roll[k])
        st.step = c["T"]

    times_gae, times_upd, launches = [], [], 0
    for i in range(reps + 1):
        load()
        torch.cuda.synchronize()
        l0 = L.b200lab_launch_count()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        st.compute_returns(roll["last_values"], c["gamma"], c["lam"])
        e[1].record()
        loss = alg.update()
        e[2].record()
        torch.cuda.synchronize()
        if i > 0:   # first iteration = warm-up (buffer allocation)
            times_gae.append(e[0].elapsed_time(e[1]))
            times_upd.append(e[1].elapsed_time(e[2]))
            launches = L.b200lab_launch_count() - l0
    gae_ms, upd_ms = min(times_gae), sorted(times_upd)[len(times_upd) // 2]
    s_iter = (gae_ms + upd_ms) * 1e-3
    return {"s_per_iter": round(s_iter, 5), "gae_ms": round(gae_ms, 4), "update_ms": round(upd_ms, 3),
            "tflops": round(ppo_flops_per_iter(c) / (upd_ms * 1e-3) / 1e12, 2), "gpu_launches": int(launches),
            "precision": {0: "fp32 FFMA (SIMT)", 1: "tf32 tcgen05 (reduced precision)",
                          2: "3xTF32 tcgen05 (fp32-accurate hi/lo split, all layers)"}[pol.precision],
            "config": f"cfg 3: N={c['N']} T={c['T']} obs={c['obs_dim']} act={c['act_dim']} hidden={c['hidden']} "
                      f"{c['epochs']} epochs x {c['mini_batches']} minibatches of {c['N'] * c['T'] // c['mini_batches']}",
            "loss": {k: round(float(v), 6) for k, v in loss.items()}}


def cpu_ppo_baseline(budget_updates=2, threads=None):
    """Oracle port of PPO.update on the host cores: `budget_updates` of the 20 minibatch updates of one iteration at
    config-3 shapes (+ GAE), scaled to s/iter."""
    from oracle import rsl_rl as O
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    c = PPO_CFG
    g = torch.Generator().manual_seed(7)
    dims = [c["obs_dim"], *c["hidden"]]
    params = {}
    for net, out in (("actor", c["act_dim"]), ("critic", 1)):
        ds = dims + [out]
        for i, (a, b) in enumerate(zip(ds[:-1], ds[1:])):
            k = 1.0 / math.sqrt(a)
            params[f"{net}.{2 * i}.weight"] = (torch.rand(b, a, generator=g) * 2 - 1) * k
            params[f"{net}.{2 * i}.bias"] = (torch.rand(b, generator=g) * 2 - 1) * k
    params["std"] = torch.ones(c["act_dim"])

    def forward(o):
        out = []
        for net in ("actor", "critic"):
            x = o
            for i in range(len(dims)):
                x = torch.nn.functional.linear(x, params[f"{net}.{2 * i}.weight"], params[f"{net}.{2 * i}.bias"])
                if i < len(dims) - 1:
                    x = torch.nn.functional.elu(x)
            out.append(x)
        return out

    roll = synthetic_rollout(c, 1234 + 3, forward)
    t0 = time.perf_counter()
    ret, adv = O.compute_returns(roll["rewards"], roll["values"], roll["dones"], roll["last_values"], c["gamma"], c["lam"])
    gae_s = time.perf_counter() - t0
    storage = dict(observations=roll["observations"], actions=roll["actions"], values=roll["values"], advantages=adv,
                   returns=ret, actions_log_prob=roll["actions_log_prob"], mu=roll["mu"], sigma=roll["sigma"])
    B = c["T"] * c["N"]
    mb = B // c["mini_batches"]
    idx = torch.randperm(B, generator=g)[: mb * budget_updates]
    cfg = dict(c, mini_batches=budget_updates, epochs=1)
    sub = {k: v.flatten(0, 1)[idx].unsqueeze(0) for k, v in storage.items()}   # [1, mb*k, .] -> k minibatches
    t0 = time.perf_counter()
    O.ppo_update(params, sub, torch.arange(mb * budget_updates), cfg, c["learning_rate"])
    upd_s = (time.perf_counter() - t0) / budget_updates * c["mini_batches"] * c["epochs"]
    return {"s_per_iter": round(gae_s + upd_s, 3), "gae_ms": round(gae_s * 1e3, 2), "cores": threads, "kind": "port",
            "sample": f"GAE + {budget_updates} of {c['mini_batches'] * c['epochs']} minibatch updates (scaled), torch CPU autograd"}


def cpu_baseline(spec, slabs, init, budget_s=15.0, threads=None, max_steps=10**9):
    """The oracle port of the reference torch path (device='cpu') on the host cores, bounded sample of the same
    workload: full-size env steps (4096 envs) until `budget_s` is used (at least 2)."""
    import copy

    from oracle.mdp_env import OracleEnv, n_u_reset, n_u_step, obs_dim
    from tests.mdp_common import init_oracle_state, slab_views_cpu
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    os.environ.setdefault("OMP_NUM_THREADS", str(threads))
    env = OracleEnv(copy.deepcopy(spec))
    env.mesh.enable_grid(0.25)
    init_oracle_state(env, init)
    N = spec["N"]
    g = torch.Generator().manual_seed(5)
    action = torch.randn(N, spec["A"], generator=g)
    n_done, t_used = 0, 0.0
    prev = slab_views_cpu(slabs, 0)[-1]
    k = 0
    while (t_used < budget_s or n_done < 2) and k < slabs["root_state_w"].shape[0] - 1 and n_done < max_steps:
        views = slab_views_cpu(slabs, k + 1)
        u = (torch.rand(N, n_u_reset(spec), generator=g), torch.rand(N, n_u_step(spec), generator=g),
             torch.rand(N, obs_dim(spec), generator=g))
        t0 = time.perf_counter()
        env.step(action, views, *u, prev)
        dt = time.perf_counter() - t0
        prev = views[-1]
        k += 1
        if k == 1:
            continue  # warm-up step (thread pools, lazy init)
        n_done += 1
        t_used += dt
    v = N * n_done / t_used
    return {"value": round(v, 1), "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{n_done} full env steps of {N} envs (after 1 warm-up), torch CPU oracle + grid-prefiltered fp64 ray cast",
            "ms_per_step": round(t_used / n_done * 1e3, 2)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg, spec, slabs, init = build_workload("cpu", seed=0)
    # each "step" of this arm = one full env step on the CPU; bounded so W+K steps end within minutes
    cb = cpu_baseline(spec, slabs, init, budget_s=20.0, max_steps=max(2, args.steps))
    out = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": workload_name(spec), "envs_per_gpu": spec["N"],
                      "arm": "reference torch path restated (oracle port) on the host cores; each step = one full env step"},
           "cpu_baseline": cb, "ppo_update": cpu_ppo_baseline(), "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    return json.dumps(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--envs", type=int, default=ENVS_PER_GPU,
                    help="envs per GPU; anything but the default is an exploration run, not the headline workload")
    args = ap.parse_args()
    if args.envs != ENVS_PER_GPU:
        globals()["ENVS_PER_GPU"] = args.envs
        globals()["K_SLABS"] = max(4, min(32, (32 * 4096) // args.envs))   # keep the cycled slabs > L2 but bounded
        PPO_CFG["N"] = args.envs
    args.warmup = max(args.warmup, 3)
    # The contract is ONE JSON line on stdout.  Native libraries write there too (NCCL prints its version banner to fd 1
    # under NCCL_DEBUG=VERSION), so fd 1 is pointed at stderr for the run and the JSON line goes to the saved descriptor.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    line = run_reference(args) if args.impl == "reference" else run_ours(args)
    sys.stdout.flush()
    if line is not None:
        os.write(real_stdout, (line + "\n").encode())


if __name__ == "__main__":
    main()
