"""Class modifiers (LAB/utils/modifiers/modifier.py:69-260) against their difference equations written out."""
import torch

from isaaclab_b200.utils import modifiers as M


def test_digital_filter_is_the_difference_equation():
    A, B = [0.5, -0.2], [0.3, 0.4, 0.1]
    f = M.DigitalFilter(M.DigitalFilterCfg(A=A, B=B), (5, 3), "cpu")
    g = torch.Generator().manual_seed(0)
    xs, ys = [torch.zeros(5, 3)] * 2, [torch.zeros(5, 3)] * 2       # x_{-1}, x_{-2}, y_{-1}, y_{-2}
    for i in range(9):
        x = torch.randn(5, 3, generator=g)
        want = B[0] * x + B[1] * xs[0] + B[2] * xs[1] - A[0] * ys[0] - A[1] * ys[1]
        got = f(x)
        torch.testing.assert_close(got, want, rtol=1e-6, atol=1e-6)
        xs, ys = [x, xs[0]], [want, ys[0]]
        if i == 4:   # reset two envs: their histories start over
            f.reset([1, 3])
            for h in (xs, ys):
                for t in h:
                    t[[1, 3]] = 0.0
    # moving average = B only
    ma = M.DigitalFilter(M.DigitalFilterCfg(A=[], B=[0.25] * 4), (2, 1), "cpu")
    out = [ma(torch.ones(2, 1)) for _ in range(5)]
    assert [round(float(o[0, 0]), 6) for o in out] == [0.25, 0.5, 0.75, 1.0, 1.0]


def test_integrator_is_trapezoidal():
    it = M.Integrator(M.IntegratorCfg(dt=0.1), (4, 2), "cpu")
    prev, total = torch.zeros(4, 2), torch.zeros(4, 2)
    g = torch.Generator().manual_seed(1)
    for i in range(6):
        x = torch.randn(4, 2, generator=g)
        total = total + (x + prev) / 2 * 0.1
        torch.testing.assert_close(it(x).clone(), total, rtol=1e-6, atol=1e-6)
        prev = x
        if i == 2:
            it.reset([0])
            prev[0], total[0] = 0.0, 0.0
