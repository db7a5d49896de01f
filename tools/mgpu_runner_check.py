"""Multi-GPU check of the PPO exchange step under torchrun (one process per GPU, NCCL):
  torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_runner_check.py
Each rank owns its own env shard (different seed => different rollouts); after `learn` the parameters, Adam moments and
the device-side learning rate and the observation-normaliser statistics must be bit-identical on every rank (ONE all-reduce of [grads | stats] per minibatch,
KL-adaptive LR recomputed identically, RSL/algorithms/ppo.py:272-291,441-469), and must differ from a rank-local update."""
import copy
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    from isaaclab_b200.envs.manager_based_rl_env import ManagerBasedRLEnv
    from isaaclab_b200.rsl_rl import OnPolicyRunner, RslRlVecEnvWrapper
    from isaaclab_b200.tasks import TASKS
    from tests.mdp_common import make_spec
    from tests.test_gpu_runner import TRAIN_CFG
    task = "Isaac-Velocity-Flat-Anymal-C-v0"
    spec = make_spec(task, 256, seed=rank)
    env = RslRlVecEnvWrapper(ManagerBasedRLEnv(TASKS[task](), spec=spec, device=f"cuda:{local}", seed=rank))
    torch.manual_seed(100 + rank)  # different initial weights per rank: broadcast_parameters must equalise them
    cfg = copy.deepcopy(TRAIN_CFG)
    cfg["policy"].update(actor_obs_normalization=True, critic_obs_normalization=True)   # global normaliser statistics
    runner = OnPolicyRunner(env, cfg, log_dir=None, device=f"cuda:{local}")
    assert runner.is_distributed and runner.alg.is_multi_gpu and runner.alg.gpu_world_size == world
    runner.learn(3)
    pol, alg = runner.alg.policy, runner.alg
    nrm = pol.actor_obs_normalizer
    # parameters, Adam moments, learning rate AND the observation-normaliser statistics (learned from the union of the
    # ranks' batches through one all-reduce per update) must be bit-identical on every rank
    blob = torch.cat([pol.flat_param, alg.exp_avg, alg.exp_avg_sq, alg._lr, nrm._mean.flatten(), nrm._var.flatten(),
                      pol.critic_obs_normalizer._mean.flatten(), nrm.count.float().reshape(1)])
    expect_count = 3 * TRAIN_CFG["num_steps_per_env"] * 256 * world
    gathered = [torch.empty_like(blob) for _ in range(world)]
    dist.all_gather(gathered, blob)
    same = all(torch.equal(gathered[0], g) for g in gathered[1:])
    finite = bool(torch.isfinite(blob).all())
    # rollouts do differ between ranks (env shards are independent)
    obs = env.get_observations()["policy"]
    osum = [torch.empty(1, device=obs.device) for _ in range(world)]
    dist.all_gather(osum, obs.sum().reshape(1))
    shards_differ = len({float(o) for o in osum}) == world
    if rank == 0:
        print(f"MGPU_CHECK world={world} params_identical={same} finite={finite} shards_differ={shards_differ} "
              f"lr={alg.learning_rate:.6g} normalizer_count={int(nrm.count)} (expected {expect_count}) "
              f"normalizer_mean_abs={float(nrm._mean.abs().mean()):.4f}")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if (same and finite and shards_differ and int(nrm.count) == expect_count) else 1)


if __name__ == "__main__":
    main()
