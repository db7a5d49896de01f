"""The reference arm of bench.py: the UNMODIFIED reference (baseline/_ref, or /root/reference where it exists) timed on
the host cores.

`time_env_step`: the reference's own `ManagerBasedRLEnv.step` -- real Action / Observation / Reward / Termination /
Command / Event / Curriculum managers built from the reference's task cfg classes, the real `Articulation`
actuator pass (TorchScript ActuatorNet*), `ContactSensor`, `RayCaster` -- over replayed PhysX tensor views
(tests/golden/gen_mdp.py builds it; the same construction generated the golden fixtures).  Closed parts only are
replaced: PhysX views -> replayed synthetic state; warp's `raycast_mesh` (warp-lang is not installable) -> the oracle's
fp64 closest-hit query with its exact xy pre-filter.
`time_ppo_update`: unmodified rsl_rl `PPO.update` / `RolloutStorage.compute_returns` on `device="cpu"`.
Not product code: only bench.py's reference / cpu_baseline legs import this."""
from __future__ import annotations

import os
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def available() -> bool:
    from tests.golden import ref_harness
    return ref_harness.available()


def time_env_step(task, spec, slabs, init, budget_s=20.0, max_steps=10**9, warmup=1, threads=None):
    """Full-size env steps of the unmodified reference until `budget_s` of CPU time is used (at least 2)."""
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    os.environ.setdefault("OMP_NUM_THREADS", str(threads))
    from oracle import build as obuild
    obuild.build()
    from tests.golden import gen_mdp   # installs the stub loader, imports the reference
    import copy
    spec = copy.deepcopy(spec)
    spec["joint_pos_limits"] = gen_mdp.hard_joint_limits(task, spec)
    gen_mdp.ref_ray_caster.raycast_mesh = gen_mdp._raycast_mesh_stand_in
    N = spec["N"]
    g = torch.Generator().manual_seed(5)
    action = torch.randn(N, spec["A"], generator=g)
    with tempfile.TemporaryDirectory() as tmp:
        env, _ = gen_mdp.build_reference_env(task, spec, {k: v.clone() for k, v in slabs.items()}, init, tmp, wrap=True)
        n_done, t_used = 0, 0.0
        with torch.inference_mode():
            for _ in range(warmup):
                env.step(action.clone())
            while (t_used < budget_s or n_done < 2) and n_done < max_steps:
                t0 = time.perf_counter()
                env.step(action.clone())
                t_used += time.perf_counter() - t0
                n_done += 1
    from tests.golden import ref_harness
    return {"value": round(N * n_done / t_used, 1), "unit": "env-steps/s", "cores": threads, "kind": "reference",
            "ms_per_step": round(t_used / n_done * 1e3, 2),
            "sample": f"{n_done} full env steps of {N} envs (after {warmup} warm-up) of the unmodified reference "
                      f"ManagerBasedRLEnv.step on CPU ({os.path.relpath(ref_harness.REFERENCE_ROOT, ROOT) if ref_harness.REFERENCE_ROOT.startswith(ROOT) else ref_harness.REFERENCE_ROOT}); "
                      "PhysX views replayed, warp ray query -> fp64 closest hit with xy pre-filter"}


def time_ppo_update(c, threads=None, reps=1):
    """Unmodified rsl_rl 3.0.0 on CPU: a T-step rollout through PPO.act / process_env_step on synthetic observations,
    then `compute_returns` (GAE) and one full `PPO.update` (epochs x minibatches) timed.  `c`: N, T, obs_dim, act_dim,
    hidden, epochs, mini_batches and the PPO hyper-parameters (bench.PPO_CFGS)."""
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    from tests.golden import ref_harness
    ref_harness.install()
    from rsl_rl.algorithms import PPO
    from rsl_rl.modules import ActorCritic
    TD = ref_harness.TensorDict
    torch.manual_seed(1234)
    N, T, D, A = c["N"], c["T"], c["obs_dim"], c["act_dim"]
    g = torch.Generator().manual_seed(1234)
    obs = TD({"policy": torch.randn(N, D, generator=g)}, batch_size=[N])
    policy = ActorCritic(obs, {"policy": ["policy"], "critic": ["policy"]}, A, actor_hidden_dims=list(c["hidden"]),
                         critic_hidden_dims=list(c["hidden"]), activation="elu", init_noise_std=1.0)
    alg = PPO(policy, num_learning_epochs=c["epochs"], num_mini_batches=c["mini_batches"], clip_param=c["clip_param"],
              gamma=c["gamma"], lam=c["lam"], value_loss_coef=c["value_loss_coef"], entropy_coef=c["entropy_coef"],
              learning_rate=c["learning_rate"], max_grad_norm=c["max_grad_norm"], use_clipped_value_loss=True,
              schedule="adaptive", desired_kl=c["desired_kl"], device="cpu")
    alg.init_storage("rl", N, T, obs, [A])
    gae_s, upd_s = [], []
    for _ in range(reps):
        for t in range(T):
            with torch.inference_mode():
                alg.act(obs)
                rewards = torch.randn(N, generator=g) * 0.05
                dones = (torch.rand(N, generator=g) < 0.01).long()
                obs = TD({"policy": torch.randn(N, D, generator=g)}, batch_size=[N])
                alg.process_env_step(obs, rewards, dones, {"time_outs": (torch.rand(N, generator=g) < 0.002) & (dones > 0)})
        t0 = time.perf_counter()
        with torch.inference_mode():
            alg.compute_returns(obs)
        t1 = time.perf_counter()
        loss = alg.update()
        t2 = time.perf_counter()
        gae_s.append(t1 - t0)
        upd_s.append(t2 - t1)
    return {"s_per_iter": round(min(gae_s) + min(upd_s), 3), "gae_ms": round(min(gae_s) * 1e3, 2),
            "update_ms": round(min(upd_s) * 1e3, 1), "cores": threads, "kind": "reference",
            "sample": f"unmodified rsl_rl 3.0.0 on CPU: GAE + full PPO.update ({c['epochs']} epochs x {c['mini_batches']} "
                      f"minibatches of {N * T // c['mini_batches']}), N={N} T={T} obs={D} hidden={list(c['hidden'])}",
            "loss": {k: round(float(v), 6) for k, v in loss.items()}}
