"""`b200lab_gae_static`: the allocation-free variant of `b200lab_gae` (per-CTA partial sums in a library-owned
`static __device__` array instead of caller scratch; include/b200lab.h).  It has no counterpart in the reference; it was
added for callers that cannot allocate.

Checked through the C ABI against the existing entry point on the same inputs, two calls in a row (the second one smaller,
so the shared array still holds the first call's partial sums beyond the slots the second call uses), within the
tolerance the GAE parity tests use (rtol 1e-5 / atol 1e-6 on returns, 2e-6 on normalised advantages) -- and bitwise, since
both entry points launch the same kernels with the same geometry.

This file is named so that pytest collects it last."""
import pytest
import torch

from isaaclab_b200 import _lib

pytestmark = pytest.mark.gpu


def _inputs(T, N, seed):
    g = torch.Generator().manual_seed(seed)
    r = (torch.randn(T, N, 1, generator=g) * 0.05).cuda()
    v = torch.randn(T, N, 1, generator=g).cuda()
    d = (torch.rand(T, N, 1, generator=g) < 0.02).to(torch.uint8).cuda()
    lv = torch.randn(N, 1, generator=g).cuda()
    return r, v, d, lv


def _gae_caller_scratch(L, r, v, d, lv, T, N, normalize):
    p = _lib.ptr
    ret, adv = torch.empty(T, N, 1, device="cuda"), torch.empty(T, N, 1, device="cuda")
    scratch = torch.zeros(L.b200lab_gae_scratch_bytes(T, N) // 8, dtype=torch.float64, device="cuda")
    _lib.check(L.b200lab_gae(p(r), p(v), p(d), p(lv), p(ret), p(adv), T, N, 0.99, 0.95, normalize, p(scratch),
                             _lib.current_stream()))
    torch.cuda.synchronize()
    return ret, adv


def test_gae_static_matches_gae_on_two_calls_in_a_row():
    L = _lib.lib()
    p = _lib.ptr
    assert _lib.consts["B200LAB_GAE_STATIC_MAX_ENVS"] == 262144
    # first call large, second small and ragged (N not a multiple of the 64-env block): stale partials stay behind
    for call, (T, N, normalize) in enumerate([(24, 4096, 1), (16, 1000, 1), (24, 4096, 0)]):
        r, v, d, lv = _inputs(T, N, seed=10 + call)
        want_ret, want_adv = _gae_caller_scratch(L, r, v, d, lv, T, N, normalize)
        ret, adv = torch.full((T, N, 1), float("nan"), device="cuda"), torch.full((T, N, 1), float("nan"), device="cuda")
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0] if call > 0 else None   # call 0 may resolve the symbol address
        _lib.check(L.b200lab_gae_static(p(r), p(v), p(d), p(lv), p(ret), p(adv), T, N, 0.99, 0.95, normalize,
                                        _lib.current_stream()))
        torch.cuda.synchronize()
        if free_before is not None:
            assert torch.cuda.mem_get_info()[0] == free_before, "b200lab_gae_static changed the device's free memory"
        torch.testing.assert_close(ret, want_ret, rtol=1e-5, atol=1e-6, msg=lambda m: f"call {call} returns: {m}")
        torch.testing.assert_close(adv, want_adv, rtol=1e-5, atol=2e-6, msg=lambda m: f"call {call} advantages: {m}")
        assert torch.equal(ret, want_ret) and torch.equal(adv, want_adv), f"call {call}: not bitwise b200lab_gae's result"


def test_gae_static_rejects_sizes_beyond_the_maximum_and_null_pointers():
    L = _lib.lib()
    p = _lib.ptr
    x = torch.zeros(8, device="cuda")
    u = torch.zeros(8, dtype=torch.uint8, device="cuda")
    too_many = _lib.consts["B200LAB_GAE_STATIC_MAX_ENVS"] + 1
    rc = L.b200lab_gae_static(p(x), p(x), p(u), p(x), p(x), p(x), 1, too_many, 0.99, 0.95, 1, _lib.current_stream())
    assert rc == _lib.consts["B200LAB_ERR_ARG"]
    msg = L.b200lab_last_error().decode()
    assert "B200LAB_GAE_STATIC_MAX_ENVS" in msg and str(too_many) in msg, msg
    assert L.b200lab_gae_static(None, p(x), p(u), p(x), p(x), p(x), 1, 8, 0.99, 0.95, 1, _lib.current_stream()) != 0
    assert L.b200lab_gae_static(p(x), p(x), p(u), p(x), p(x), p(x), 0, 8, 0.99, 0.95, 1, _lib.current_stream()) != 0
    torch.cuda.synchronize()   # nothing was launched by the rejected calls; the context is healthy
