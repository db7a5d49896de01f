"""Quick numerical check of the tcgen05 GEMM paths against torch fp64 (run on the GPU box under `timeout`)."""
import sys
import time

import torch

sys.path.insert(0, ".")
from isaaclab_b200 import _lib  # noqa: E402

L = _lib.lib()


def run(M, K, N, precision, elu=1):
    g = torch.Generator().manual_seed(M + K + N)
    X, W, b = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / K ** 0.5, torch.randn(N, generator=g)
    dY = torch.randn(M, N, generator=g) / M
    Xr, Wr, br = X.double().requires_grad_(), W.double().requires_grad_(), b.double().requires_grad_()
    Yr = torch.nn.functional.linear(Xr, Wr, br)
    if elu:
        Yr = torch.nn.functional.elu(Yr)
    Yr.backward(dY.double())
    Xc, Wc, bc, dYc = X.cuda(), W.cuda(), b.cuda(), dY.cuda()
    Y = torch.empty(M, N, device="cuda")
    s = _lib.current_stream()
    _lib.check(L.b200lab_linear_fwd(_lib.ptr(Xc), _lib.ptr(Wc), _lib.ptr(bc), _lib.ptr(Y), M, K, N, elu, precision, s))
    torch.cuda.synchronize()
    e_fwd = ((Y.cpu().double() - Yr.detach()).abs().max() / Yr.detach().abs().max()).item()
    dZ, dX = torch.empty(M, N, device="cuda"), torch.empty(M, K, device="cuda")
    dW, db = torch.empty(N, K, device="cuda"), torch.empty(N, device="cuda")
    ws = torch.empty(L.b200lab_linear_bwd_workspace_bytes(M, K, N) // 4 + 1, device="cuda")
    _lib.check(L.b200lab_linear_bwd(_lib.ptr(Xc), _lib.ptr(Wc), _lib.ptr(Y), _lib.ptr(dYc), _lib.ptr(dZ), _lib.ptr(dX),
                                    _lib.ptr(dW), _lib.ptr(db), M, K, N, elu, precision, _lib.ptr(ws), s))
    torch.cuda.synchronize()
    e_dx = ((dX.cpu().double() - Xr.grad).abs().max() / Xr.grad.abs().max()).item()
    e_dw = ((dW.cpu().double() - Wr.grad).abs().max() / Wr.grad.abs().max()).item()
    # timing
    for _ in range(3):
        L.b200lab_linear_fwd(_lib.ptr(Xc), _lib.ptr(Wc), _lib.ptr(bc), _lib.ptr(Y), M, K, N, elu, precision, s)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        L.b200lab_linear_fwd(_lib.ptr(Xc), _lib.ptr(Wc), _lib.ptr(bc), _lib.ptr(Y), M, K, N, elu, precision, s)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 100
    print(f"M={M} K={K} N={N} precision={precision}: rel err fwd {e_fwd:.2e} dX {e_dx:.2e} dW {e_dw:.2e}  "
          f"fwd {us:.1f} us = {2 * M * K * N / us / 1e6:.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    shapes = [(256, 64, 64), (384, 96, 128), (1000, 128, 256), (24576, 235, 512), (24576, 512, 256), (24576, 256, 128)]
    for p in (0, 1, 2):
        for (M, K, N) in shapes:
            run(M, K, N, p)
