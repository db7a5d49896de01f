"""GPU parity of the rsl_rl hot path (GAE, storage, minibatch gather, MLP fwd/bwd, PPO update) through the C ABI,
against the golden fixtures produced by the unmodified reference and against the CPU oracle on seeded inputs."""
import os

import pytest
import torch

from oracle import rsl_rl as O

pytestmark = pytest.mark.gpu


def _load(golden_dir, name):
    return torch.load(os.path.join(golden_dir, f"rsl_rl_{name}.pt"), weights_only=False)


def _mk(fx, dev="cuda"):
    from isaaclab_b200.rsl_rl import PPO, ActorCritic, TensorDict
    c = fx["cfg"]
    obs0 = TensorDict({"policy": torch.zeros(c["N"], c["obs_dim"], device=dev)}, batch_size=[c["N"]])
    pol = ActorCritic(obs0, {"policy": ["policy"], "critic": ["policy"]}, c["act_dim"], actor_hidden_dims=c["hidden"],
                      critic_hidden_dims=c["hidden"], device=dev, actor_obs_normalization=c.get("normalization", False),
                      critic_obs_normalization=c.get("normalization", False))
    pol.load_state_dict(fx["state_dict_before"])
    alg = PPO(pol, num_learning_epochs=c["epochs"], num_mini_batches=c["mini_batches"], clip_param=c["clip_param"],
              gamma=c["gamma"], lam=c["lam"], value_loss_coef=c["value_loss_coef"], entropy_coef=c["entropy_coef"],
              learning_rate=c["learning_rate"], max_grad_norm=c["max_grad_norm"], desired_kl=c["desired_kl"], device=dev)
    alg.init_storage("rl", c["N"], c["T"], obs0, [c["act_dim"]])
    return pol, alg, TensorDict


@pytest.mark.parametrize("name", ["small", "ragged", "norm"])
def test_rollout_gae_update_vs_reference_fixture(golden_dir, name):
    fx = _load(golden_dir, name)
    c = fx["cfg"]
    pol, alg, TensorDict = _mk(fx)
    dev = "cuda"
    for t, s in enumerate(fx["steps"]):
        obs = TensorDict({"policy": s["obs"].to(dev)}, batch_size=[c["N"]])
        eps = ((s["actions"] - s["mu"]) / s["sigma"]).to(dev)
        a = alg.act(obs, eps=eps)
        torch.testing.assert_close(pol.action_mean.cpu(), s["mu"], rtol=1e-5, atol=2e-6)
        torch.testing.assert_close(alg.transition.values.cpu(), s["values"], rtol=1e-5, atol=2e-6)
        torch.testing.assert_close(a.cpu(), s["actions"], rtol=1e-5, atol=2e-6)
        torch.testing.assert_close(alg.transition.actions_log_prob.cpu(), s["log_prob"], rtol=1e-5, atol=2e-5)
        extras = {} if s["time_outs"] is None else {"time_outs": s["time_outs"].to(dev)}
        nxt = fx["steps"][t + 1]["obs"] if t + 1 < len(fx["steps"]) else fx["last_obs"]   # what update_normalization learns
        alg.process_env_step(TensorDict({"policy": nxt.to(dev)}, batch_size=[c["N"]]), s["rewards"].to(dev),
                             s["dones"].to(dev), extras)
    st = alg.storage
    torch.testing.assert_close(st.rewards.cpu(), fx["storage"]["rewards"], rtol=1e-5, atol=1e-6)
    assert torch.equal(st.dones.cpu(), fx["storage"]["dones"])
    assert torch.equal(st.observations["policy"].cpu(), fx["storage"]["observations"])
    # load the exact reference rollout so GAE / update parity is not blurred by MLP rounding
    for k in ("actions", "rewards", "values", "actions_log_prob", "mu", "sigma"):
        getattr(st, k).copy_(fx["storage"][k])
    st.compute_returns(fx["last_values"].to(dev), c["gamma"], c["lam"])
    torch.testing.assert_close(st.returns.cpu(), fx["storage"]["returns"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(st.advantages.cpu(), fx["storage"]["advantages"], rtol=1e-5, atol=1e-6)
    loss = alg.update(indices=fx["indices"].to(dev))
    for k, v in fx["state_dict_after"].items():
        got = pol.state_dict()[k].cpu()
        # Adam divides by sqrt(v)+eps: where |grad| ~ 1e-8 the update is sign-like and amplifies 1e-9 GEMM rounding
        # differences, so a handful of elements may move by up to ~lr; everything else must agree tightly.
        torch.testing.assert_close(got, v, rtol=2e-4, atol=5e-4, msg=lambda m, k=k: f"{k}: {m}")
        tight = ((got - v).abs() <= 5e-6 + 2e-4 * v.abs()).float().mean().item()
        assert tight >= 0.995, f"{k}: only {tight:.4f} of the elements within rtol 2e-4 / atol 5e-6"
    for k in ("value_function", "surrogate", "entropy"):
        assert loss[k] == pytest.approx(fx["loss_dict"][k], rel=1e-4, abs=1e-6)
    assert alg.learning_rate == pytest.approx(float(fx["lr_per_update"][-1]), rel=1e-6)


def test_gae_vs_oracle_full_size_and_edges():
    from isaaclab_b200 import _lib
    L = _lib.lib()
    for (T, N, pd) in [(24, 4096, 0.01), (1, 1, 0.5), (3, 65, 1.0), (24, 16384, 0.0), (7, 1000, 0.3)]:
        g = torch.Generator().manual_seed(T * 100003 + N)
        r = torch.randn(T, N, 1, generator=g) * 0.05
        v = torch.randn(T, N, 1, generator=g)
        d = (torch.rand(T, N, 1, generator=g) < pd).to(torch.uint8)
        lv = torch.randn(N, 1, generator=g)
        for norm in (True, False):
            if T * N < 2 and norm:
                continue
            ret_o, adv_o = O.compute_returns(r, v, d, lv, 0.99, 0.95, norm)
            ret = torch.empty(T, N, 1, device="cuda")
            adv = torch.empty(T, N, 1, device="cuda")
            scratch = torch.zeros(L.b200lab_gae_scratch_bytes(T, N) // 8, dtype=torch.float64, device="cuda")
            rc, vc, dc, lvc = r.cuda(), v.cuda(), d.cuda(), lv.cuda()  # keep alive: raw pointers are borrowed
            _lib.check(L.b200lab_gae(_lib.ptr(rc), _lib.ptr(vc), _lib.ptr(dc), _lib.ptr(lvc),
                                     _lib.ptr(ret), _lib.ptr(adv), T, N, 0.99, 0.95, int(norm), _lib.ptr(scratch),
                                     _lib.current_stream()))
            torch.testing.assert_close(ret.cpu(), ret_o, rtol=1e-5, atol=1e-6)
            torch.testing.assert_close(adv.cpu(), adv_o, rtol=1e-5, atol=2e-6)


def test_gather_is_bit_exact_and_is_permutation_invariant():
    from isaaclab_b200.rsl_rl import RolloutStorage, TensorDict
    T, N, D, A = 6, 50, 37, 5
    obs = TensorDict({"policy": torch.zeros(N, D, device="cuda")}, batch_size=[N])
    st = RolloutStorage("rl", N, T, obs, [A], "cuda")
    g = torch.Generator().manual_seed(3)
    for k in ("actions", "rewards", "values", "actions_log_prob", "mu", "sigma", "returns", "advantages"):
        getattr(st, k).copy_(torch.randn(getattr(st, k).shape, generator=g))
    st.observations["policy"].copy_(torch.randn(T, N, D, generator=g))
    idx = torch.randperm(T * N, generator=g)
    seen = []
    for i, mb in enumerate(st.mini_batch_generator(3, 2, indices=idx.cuda())):
        sl = idx[(i % 3) * 100:(i % 3 + 1) * 100]
        assert torch.equal(mb[0]["policy"].cpu(), st.observations["policy"].flatten(0, 1).cpu()[sl])
        assert torch.equal(mb[1].cpu(), st.actions.flatten(0, 1).cpu()[sl])
        assert torch.equal(mb[2].cpu(), st.values.flatten(0, 1).cpu()[sl])
        assert torch.equal(mb[3].cpu(), st.advantages.flatten(0, 1).cpu()[sl])
        assert torch.equal(mb[4].cpu(), st.returns.flatten(0, 1).cpu()[sl])
        assert torch.equal(mb[5].cpu(), st.actions_log_prob.flatten(0, 1).cpu()[sl])
        assert torch.equal(mb[6].cpu(), st.mu.flatten(0, 1).cpu()[sl])
        assert torch.equal(mb[7].cpu(), st.sigma.flatten(0, 1).cpu()[sl])
        seen.append(mb[2].sum().item())
    assert seen[:3] == seen[3:]  # same permutation reused by every epoch


@pytest.mark.parametrize("M,K,N", [(24576, 235, 512), (4096, 48, 128), (1000, 128, 12), (333, 77, 1), (64, 512, 256)])
def test_linear_fwd_bwd_vs_torch_fp32(M, K, N):
    from isaaclab_b200 import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(M + K + N)
    X = torch.randn(M, K, generator=g)
    W = torch.randn(N, K, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g)
    dY = torch.randn(M, N, generator=g) / M
    Xr, Wr, br = X.double().requires_grad_(), W.double().requires_grad_(), b.double().requires_grad_()
    Yr = torch.nn.functional.elu(torch.nn.functional.linear(Xr, Wr, br))
    Yr.backward(dY.double())
    Xc, Wc, bc, dYc = X.cuda(), W.cuda(), b.cuda(), dY.cuda()
    Y = torch.empty(M, N, device="cuda")
    s = _lib.current_stream()
    _lib.check(L.b200lab_linear_fwd(_lib.ptr(Xc), _lib.ptr(Wc), _lib.ptr(bc), _lib.ptr(Y), M, K, N, 1, 0, s))
    torch.testing.assert_close(Y.cpu().double(), Yr.detach(), rtol=1e-5, atol=1e-5)
    dZ = torch.empty(M, N, device="cuda"); dX = torch.empty(M, K, device="cuda")
    dW = torch.empty(N, K, device="cuda"); db = torch.empty(N, device="cuda")
    ws = torch.empty(L.b200lab_linear_bwd_workspace_bytes(M, K, N) // 4 + 1, device="cuda")
    _lib.check(L.b200lab_linear_bwd(_lib.ptr(Xc), _lib.ptr(Wc), _lib.ptr(Y), _lib.ptr(dYc), _lib.ptr(dZ), _lib.ptr(dX),
                                    _lib.ptr(dW), _lib.ptr(db), M, K, N, 1, 0, _lib.ptr(ws), s))
    torch.testing.assert_close(dX.cpu().double(), Xr.grad, rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(dW.cpu().double(), Wr.grad, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(db.cpu().double(), br.grad, rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("precision,tol", [(2, 3e-5), (1, 3e-3)], ids=["tf32x3", "tf32"])
@pytest.mark.parametrize("M,K,N", [(24576, 512, 256), (24576, 235, 512), (1000, 128, 256), (384, 96, 128), (333, 77, 40),
                                   (4096, 256, 128), (24576, 128, 12), (5000, 128, 1), (777, 4, 32), (130, 33, 19),
                                   (20011, 236, 300), (9999, 512, 303), (40000, 64, 640)])
def test_linear_tcgen05_vs_fp64(M, K, N, precision, tol):
    """tcgen05 kind::tf32 GEMM paths (csrc/mlp_tc.cu): forward with fused bias+ELU, dX and split-K dW, against torch fp64.
    3xTF32 (precision 2) must stay at fp32-level accuracy; single TF32 (precision 1) at TF32 level.  Errors are measured
    relative to the largest magnitude of each result (the layers' outputs are O(1), the gradients O(1/M))."""
    from isaaclab_b200 import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(M + K + N)
    X = torch.randn(M, K, generator=g)
    W = torch.randn(N, K, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g)
    dY = torch.randn(M, N, generator=g) / M
    Xr, Wr, br = X.double().requires_grad_(), W.double().requires_grad_(), b.double().requires_grad_()
    Yr = torch.nn.functional.elu(torch.nn.functional.linear(Xr, Wr, br))
    Yr.backward(dY.double())
    Xc, Wc, bc, dYc = X.cuda(), W.cuda(), b.cuda(), dY.cuda()
    Y = torch.empty(M, N, device="cuda")
    s = _lib.current_stream()
    _lib.check(L.b200lab_linear_fwd(_lib.ptr(Xc), _lib.ptr(Wc), _lib.ptr(bc), _lib.ptr(Y), M, K, N, 1, precision, s))
    dZ = torch.empty(M, N, device="cuda"); dX = torch.empty(M, K, device="cuda")
    dW = torch.empty(N, K, device="cuda"); db = torch.empty(N, device="cuda")
    ws = torch.empty(L.b200lab_linear_bwd_workspace_bytes(M, K, N) // 4 + 1, device="cuda")
    _lib.check(L.b200lab_linear_bwd(_lib.ptr(Xc), _lib.ptr(Wc), _lib.ptr(Y), _lib.ptr(dYc), _lib.ptr(dZ), _lib.ptr(dX),
                                    _lib.ptr(dW), _lib.ptr(db), M, K, N, 1, precision, _lib.ptr(ws), s))
    torch.cuda.synchronize()

    def rel(a, r):
        return ((a.cpu().double() - r).abs().max() / r.abs().max()).item()
    assert rel(Y, Yr.detach()) < tol, f"fwd {rel(Y, Yr.detach()):.2e}"
    assert rel(dX, Xr.grad) < tol, f"dX {rel(dX, Xr.grad):.2e}"
    assert rel(dW, Wr.grad) < tol, f"dW {rel(dW, Wr.grad):.2e}"
    assert rel(db, br.grad) < tol, f"db {rel(db, br.grad):.2e}"   # through ELU'(Y): inherits the forward error


@pytest.mark.parametrize("precision,tol", [(2, 3e-5), (1, 3e-3), (0, 1e-5)], ids=["tf32x3", "tf32", "ffma"])
@pytest.mark.parametrize("M,dims", [(24576, [236, 512, 256, 128, 12]), (1000, [48, 128, 130, 1]), (333, [77, 64, 32, 5])])
def test_linear_bwd_chain_vs_fp64(M, dims, precision, tol):
    """b200lab_linear_bwd_chain over a Linear+ELU ... Linear stack (RSL/networks/mlp.py:60-80): every layer below the last
    receives its dZ and bias gradient from the call of the layer above (ELU_IN there, DZ_GIVEN here); all weight, bias
    and input gradients against torch autograd in fp64."""
    from isaaclab_b200 import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(M + sum(dims))
    n = len(dims) - 1
    X = torch.randn(M, dims[0], generator=g)
    Ws = [torch.randn(dims[i + 1], dims[i], generator=g) / dims[i] ** 0.5 for i in range(n)]
    bs = [torch.randn(dims[i + 1], generator=g) * 0.1 for i in range(n)]
    dOut = torch.randn(M, dims[-1], generator=g) / M
    Xr = X.double().requires_grad_()
    Wr = [w.double().requires_grad_() for w in Ws]
    br = [b.double().requires_grad_() for b in bs]
    h = Xr
    for i in range(n):
        h = torch.nn.functional.linear(h, Wr[i], br[i])
        if i < n - 1:
            h = torch.nn.functional.elu(h)
    h.backward(dOut.double())
    s = _lib.current_stream()
    Xc, Wc, bc = X.cuda(), [w.cuda() for w in Ws], [b.cuda() for b in bs]
    acts, inp = [], Xc
    for i in range(n):
        Y = torch.empty(M, dims[i + 1], device="cuda")
        _lib.check(L.b200lab_linear_fwd(_lib.ptr(inp), _lib.ptr(Wc[i]), _lib.ptr(bc[i]), _lib.ptr(Y), M, dims[i], dims[i + 1],
                                        0 if i == n - 1 else 1, precision, s))
        acts.append(Y)
        inp = Y
    dW = [torch.empty_like(w) for w in Wc]
    db = [torch.empty_like(b) for b in bc]
    grads = [torch.empty(M, dims[i], device="cuda") for i in range(n)]   # grads[i] = gradient w.r.t. the input of layer i
    ws = torch.empty(max(L.b200lab_linear_bwd_workspace_bytes(M, dims[i], dims[i + 1]) for i in range(n)) // 4 + 1, device="cuda")
    dy = dOut.cuda()
    for i in reversed(range(n)):
        flags = (0 if i == n - 1 else 4) | (2 if i > 0 else 0)
        x_in = Xc if i == 0 else acts[i - 1]
        _lib.check(L.b200lab_linear_bwd_chain(_lib.ptr(x_in), _lib.ptr(Wc[i]), None, _lib.ptr(dy), None, _lib.ptr(grads[i]),
                                              _lib.ptr(dW[i]), _lib.ptr(db[i]), _lib.ptr(db[i - 1]) if i > 0 else None, M,
                                              dims[i], dims[i + 1], flags, precision, _lib.ptr(ws), s))
        dy = grads[i]
    torch.cuda.synchronize()

    def rel(a, r):
        return ((a.cpu().double() - r).abs().max() / r.abs().max()).item()
    assert rel(grads[0], Xr.grad) < tol, f"dX {rel(grads[0], Xr.grad):.2e}"
    for i in range(n):
        assert rel(dW[i], Wr[i].grad) < tol, f"dW[{i}] {rel(dW[i], Wr[i].grad):.2e}"
        assert rel(db[i], br[i].grad) < tol, f"db[{i}] {rel(db[i], br[i].grad):.2e}"
    # flags that contradict each other are refused
    rc = L.b200lab_linear_bwd_chain(_lib.ptr(Xc), _lib.ptr(Wc[0]), None, _lib.ptr(dy), None, None, _lib.ptr(dW[0]), _lib.ptr(db[0]),
                                    None, M, dims[0], dims[1], 1 | 4, precision, _lib.ptr(ws), s)
    assert rc != 0


@pytest.mark.parametrize("precision", [2, 1])
def test_linear_tcgen05_is_bitwise_repeatable(precision):
    """The persistent GEMM hands stages between a TMA warp, sweepers, the MMA lane and epilogue warps through mbarriers; a
    missing wait shows up as run-to-run differences.  30 launches of forward, dX and split-K dW on a many-tile shape with
    tails, interleaved with a kernel that dirties L2, must all be bit-identical."""
    from isaaclab_b200 import _lib
    L = _lib.lib()
    M, K, N = 20011, 236, 300
    g = torch.Generator().manual_seed(7)
    X = torch.randn(M, K, generator=g).cuda()
    W = (torch.randn(N, K, generator=g) / K ** 0.5).cuda()
    b = torch.randn(N, generator=g).cuda()
    dY = (torch.randn(M, N, generator=g) / M).cuda()
    s = _lib.current_stream()
    ws = torch.empty(L.b200lab_linear_bwd_workspace_bytes(M, K, N) // 4 + 1, device="cuda")
    junk = torch.empty(64 << 20, device="cuda")
    ref = None
    for it in range(30):
        Y = torch.full((M, N), float("nan"), device="cuda")
        dZ = torch.empty(M, N, device="cuda"); dX = torch.full((M, K), float("nan"), device="cuda")
        dW = torch.full((N, K), float("nan"), device="cuda"); db = torch.empty(N, device="cuda")
        _lib.check(L.b200lab_linear_fwd(_lib.ptr(X), _lib.ptr(W), _lib.ptr(b), _lib.ptr(Y), M, K, N, 1, precision, s))
        _lib.check(L.b200lab_linear_bwd(_lib.ptr(X), _lib.ptr(W), _lib.ptr(Y), _lib.ptr(dY), _lib.ptr(dZ), _lib.ptr(dX),
                                        _lib.ptr(dW), _lib.ptr(db), M, K, N, 1, precision, _lib.ptr(ws), s))
        junk.add_(1.0)
        out = [t.clone() for t in (Y, dX, dW, db)]
        if ref is None:
            ref = out
            assert all(torch.isfinite(t).all() for t in ref)
        else:
            for a, r, name in zip(out, ref, ("Y", "dX", "dW", "db")):
                assert torch.equal(a, r), f"{name} differs in launch {it}"


def test_linear_rejects_unknown_precision():
    from isaaclab_b200 import _lib
    L = _lib.lib()
    x = torch.zeros(64, 64, device="cuda")
    rc = L.b200lab_linear_fwd(_lib.ptr(x), _lib.ptr(x), _lib.ptr(x[0]), _lib.ptr(x.clone()), 64, 64, 64, 0, 7, _lib.current_stream())
    assert rc != 0 and b"precision" in L.b200lab_last_error()


def test_global_normalizer_statistics_equal_the_update_on_the_union_of_the_ranks_batches():
    """N1 (north_star): partial sums per rank + ONE all-reduce + merge == the reference's running update
    (normalization.py:50-66) applied to the concatenation of the ranks' batches.  Two 'ranks' are emulated on one GPU by
    adding their partial buffers (what the SUM all-reduce does); tests/test_gpu_multi.py runs the real NCCL exchange."""
    from isaaclab_b200 import _lib
    L, p, s = _lib.lib(), _lib.ptr, _lib.current_stream
    D = 235
    g = torch.Generator().manual_seed(4)
    scale, shift = 0.5 + 3.0 * torch.rand(D, generator=g), 2.0 * torch.randn(D, generator=g)
    mean_g, var_g, std_g = (torch.zeros(D, device="cuda"), torch.ones(D, device="cuda"), torch.ones(D, device="cuda"))
    mean_r, var_r, count = torch.zeros(D, dtype=torch.float64), torch.ones(D, dtype=torch.float64), 0
    for it in range(3):
        xs = [(torch.randn(m, D, generator=g) * scale + shift + it).cuda() for m in (4096, 1000)]   # unequal shards
        bufs = [torch.zeros(2 * D + 1, dtype=torch.float64, device="cuda") for _ in xs]
        for x, b in zip(xs, bufs):
            _lib.check(L.b200lab_empirical_norm_partial(p(x), x.shape[0], D, p(b), s()))
        total = bufs[0] + bufs[1]
        _lib.check(L.b200lab_empirical_norm_merge(p(total), D, count, p(mean_g), p(var_g), p(std_g), s()))
        # reference update on the union, in fp64
        x = torch.cat(xs).cpu().double()
        count += x.shape[0]
        rate = x.shape[0] / count
        mx, vx = x.mean(0), x.var(0, unbiased=False)
        delta = mx - mean_r
        mean_r = mean_r + rate * delta
        var_r = var_r + rate * (vx - var_r + delta * (mx - mean_r))
        torch.cuda.synchronize()
        assert float(total[2 * D]) == 5096.0
        torch.testing.assert_close(mean_g.cpu().double(), mean_r, rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(var_g.cpu().double(), var_r, rtol=1e-4, atol=1e-5)
        torch.testing.assert_close(std_g.cpu().double(), var_r.sqrt(), rtol=1e-4, atol=1e-5)
    # one rank: the global path equals the rank-local kernel
    x = (torch.randn(2048, D, generator=g) * scale + shift).cuda()
    m1, v1, s1 = torch.zeros(D, device="cuda"), torch.ones(D, device="cuda"), torch.ones(D, device="cuda")
    m2, v2, s2 = m1.clone(), v1.clone(), s1.clone()
    _lib.check(L.b200lab_empirical_norm_update(p(x), 2048, D, 2048, p(m1), p(v1), p(s1), s()))
    b = torch.zeros(2 * D + 1, dtype=torch.float64, device="cuda")
    _lib.check(L.b200lab_empirical_norm_partial(p(x), 2048, D, p(b), s()))
    _lib.check(L.b200lab_empirical_norm_merge(p(b), D, 0, p(m2), p(v2), p(s2), s()))
    torch.testing.assert_close(m2, m1, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(v2, v1, rtol=1e-5, atol=1e-5)
