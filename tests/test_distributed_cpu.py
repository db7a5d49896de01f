"""N>1 host path on CPU (gloo, world_size 2): the ONE all-reduce of the flat [grads | stats] buffer that replaces the
reference's per-parameter all-reduce + KL all-reduce + LR broadcast (RSL/algorithms/ppo.py:276-291,444-469), the
parameter broadcast, and env sharding.  The arithmetic after the reduce is checked against the oracle's single-process
statement of the reference semantics (mean of per-rank gradients, mean KL over ranks)."""
import os
import socket
import types

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import rsl_rl as O


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_inputs(rank, n_param):
    g = torch.Generator().manual_seed(100 + rank)
    grads = torch.randn(n_param, generator=g)
    stats = torch.rand(8, generator=g)   # per-rank sums: value loss, surrogate, entropy, kl, count, ...
    return grads, stats


def _worker(rank, world, port, n_param, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from isaaclab_b200.rsl_rl.ppo import PPO
        grads, stats = _rank_inputs(rank, n_param)
        flat_param = torch.full((n_param,), float(rank + 1))
        alg = PPO.__new__(PPO)
        alg.policy = types.SimpleNamespace(flat_param=flat_param, flat_grad_ext=torch.cat([grads, stats]))
        alg.is_multi_gpu, alg.gpu_global_rank, alg.gpu_world_size = True, rank, world
        alg.broadcast_parameters()
        alg.reduce_parameters()
        torch.save(dict(param=alg.policy.flat_param, ext=alg.policy.flat_grad_ext), os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_flat_allreduce_and_broadcast_world2(tmp_path):
    world, n_param = 2, 1000
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n_param, str(tmp_path)), nprocs=world, join=True)
    outs = [torch.load(tmp_path / f"r{r}.pt") for r in range(world)]
    # broadcast: every rank holds rank 0's parameters
    for o in outs:
        assert torch.equal(o["param"], torch.full((n_param,), 1.0))
    # reduce: identical sums on both ranks
    assert torch.equal(outs[0]["ext"], outs[1]["ext"])
    ins = [_rank_inputs(r, n_param) for r in range(world)]
    want_grad_sum = ins[0][0] + ins[1][0]
    torch.testing.assert_close(outs[0]["ext"][:n_param], want_grad_sum, rtol=0, atol=0)
    torch.testing.assert_close(outs[0]["ext"][n_param:], ins[0][1] + ins[1][1], rtol=0, atol=0)
    # reference semantics (ppo.py:459-469): every parameter's gradient becomes the MEAN over ranks; the 1/world factor
    # the CUDA grad-norm/Adam kernels fold in must reproduce the oracle's clip + Adam on the averaged gradient
    mean_grad = want_grad_sum / world
    folded_grad = outs[0]["ext"][:n_param] * (1.0 / world)
    res = []
    for gsrc in (mean_grad, folded_grad):
        params = {"w": torch.full((n_param,), 1.0)}
        grads = {"w": gsrc.clone()}
        O.clip_grad_norm(grads, 1.0)
        adam = O.Adam(params)
        adam.step(params, grads, lr=1e-3)
        res.append(params["w"])
    torch.testing.assert_close(res[0], res[1], rtol=1e-7, atol=1e-9)
    # KL for the adaptive LR: mean over ranks of the per-rank mean (ppo.py:276-279) == summed stats / world
    kl_per_rank = [float(i[1][3]) for i in ins]
    assert abs(sum(kl_per_rank) / world - float(outs[0]["ext"][n_param + 3]) / world) < 1e-7


def test_env_sharding_is_disjoint_and_seeded_per_rank():
    """Weak scaling: every rank owns its own envs (no data-path collective); distinct seeds give distinct streams."""
    from isaaclab_b200.sim.synthetic import initial_env_state
    from tests.mdp_common import make_spec
    spec = make_spec("Isaac-Velocity-Flat-Anymal-C-v0", 32)
    a, b = initial_env_state(spec, seed=7 + 0), initial_env_state(spec, seed=7 + 1)
    assert not torch.equal(a["episode_length_buf"], b["episode_length_buf"])
