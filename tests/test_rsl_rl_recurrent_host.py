"""Host logic of the recurrent policy path (CPU): trajectory splitting / padding and the recurrent minibatch generator
against what the UNMODIFIED rsl_rl produced (tests/golden/gen_rsl_rl.py --recurrent; RSL/utils/utils.py:78-141,
RSL/storage/rollout_storage.py:206-260)."""
import os

import pytest
import torch

from isaaclab_b200.rsl_rl import RolloutStorage, TensorDict, split_and_pad_trajectories, unpad_trajectories

RECURRENT = ["lstm", "gru2", "lstm2"]


def _naive_split(tensor, dones):
    """Per-env Python loop: the definition of the operation."""
    T, N = dones.shape[:2]
    trajs = []
    for n in range(N):
        cur = []
        for t in range(T):
            cur.append(tensor[t, n])
            if dones[t, n] or t == T - 1:
                trajs.append(torch.stack(cur))
                cur = []
    padded = torch.zeros(T, len(trajs), *tensor.shape[2:])
    masks = torch.zeros(T, len(trajs), dtype=torch.bool)
    for i, tr in enumerate(trajs):
        padded[:len(tr), i] = tr
        masks[:len(tr), i] = True
    return padded, masks


def test_split_and_pad_docstring_example():
    # [[a1 a2 a3 a4 | a5 a6], [b1 b2 | b3 b4 b5 | b6]]  (utils.py:84-94)
    x = torch.tensor([[1., 2, 3, 4, 5, 6], [11, 12, 13, 14, 15, 16]]).t().unsqueeze(-1)   # [T=6, N=2, 1]
    d = torch.tensor([[0, 0, 0, 1, 0, 0], [0, 1, 0, 0, 1, 0]]).t().unsqueeze(-1)
    padded, masks = split_and_pad_trajectories(x, d)
    assert padded.shape == (6, 5, 1)
    assert padded[:, :, 0].t().tolist() == [[1, 2, 3, 4, 0, 0], [5, 6, 0, 0, 0, 0], [11, 12, 0, 0, 0, 0], [13, 14, 15, 0, 0, 0],
                                            [16, 0, 0, 0, 0, 0]]
    assert masks.t().sum(1).tolist() == [4, 2, 2, 3, 1]
    back = unpad_trajectories(padded, masks)
    assert torch.equal(back, x)


@pytest.mark.parametrize("T,N,pd", [(1, 1, 0.0), (8, 5, 0.0), (8, 5, 1.0), (24, 33, 0.1), (7, 64, 0.4)])
def test_split_and_pad_matches_the_definition(T, N, pd):
    g = torch.Generator().manual_seed(T * 131 + N)
    x = torch.randn(T, N, 3, generator=g)
    d = (torch.rand(T, N, 1, generator=g) < pd)
    padded, masks = split_and_pad_trajectories(x, d)
    ref_p, ref_m = _naive_split(x, d.squeeze(-1))
    assert torch.equal(padded, ref_p) and torch.equal(masks, ref_m)
    assert torch.equal(unpad_trajectories(padded, masks), x)
    td = TensorDict({"policy": x, "critic": x[..., :2].contiguous()}, batch_size=[T, N])
    ptd, m2 = split_and_pad_trajectories(td, d.to(torch.uint8))
    assert torch.equal(ptd["policy"], ref_p) and torch.equal(ptd["critic"], ref_p[..., :2]) and torch.equal(m2, ref_m)


@pytest.mark.parametrize("name", RECURRENT)
def test_recurrent_generator_matches_reference(golden_dir, name):
    fx = torch.load(os.path.join(golden_dir, f"rsl_rl_{name}.pt"), weights_only=False)
    c, s = fx["cfg"], fx["storage"]
    obs0 = TensorDict({"policy": torch.zeros(c["N"], c["obs_dim"])}, batch_size=[c["N"]])
    st = RolloutStorage("rl", c["N"], c["T"], obs0, [c["act_dim"]], "cpu")
    st.observations["policy"].copy_(s["observations"])
    for k in ("actions", "rewards", "values", "actions_log_prob", "mu", "sigma", "returns", "advantages"):
        getattr(st, k).copy_(s[k])
    st.dones.copy_(s["dones"])
    st.saved_hidden_states_a = [h.clone() for h in s["saved_hidden_states_a"]]
    st.saved_hidden_states_c = [h.clone() for h in s["saved_hidden_states_c"]]
    got = list(st.recurrent_mini_batch_generator(c["mini_batches"], 2))
    assert len(got) == 2 * c["mini_batches"]
    for i, mb in enumerate(got):
        ref = fx["recurrent_batches"][i % c["mini_batches"]]
        assert torch.equal(mb[0]["policy"], ref["obs"]) and torch.equal(mb[9], ref["masks"])
        for j, k in enumerate(("actions", "values", "advantages", "returns", "log_prob", "mu", "sigma"), start=1):
            assert torch.equal(mb[j], ref[k]), k
        for mine, theirs in ((mb[8][0], ref["hid_a"]), (mb[8][1], ref["hid_c"])):
            if isinstance(theirs, list):
                assert len(mine) == len(theirs) and all(torch.equal(a, b) for a, b in zip(mine, theirs))
            else:
                assert torch.equal(mine, theirs)
