"""CPU tests of the oracle itself: the two independent restatements agree, the gradients are true adjoints, and the
known answers of SURVEY.md section 4 hold.  (Pinning against the reference's own kernels: test_reference_pin.py and
test_golden.py.)"""
import numpy as np
import pytest
import torch

from oracle import oracle as o
from conftest import assert_close

CASES = [
    # N  H   W   C  k  md s1 s2 pad
    (2, 9, 11, 8, 1, 4, 1, 2, 4),      # FlowNetC-like (pad == md, s2 = 2)
    (1, 10, 12, 5, 3, 3, 2, 2, 3),     # k = 3, s1 = 2, odd C
    (1, 8, 9, 4, 1, 3, 1, 1, 1),       # pad < md: output smaller than input
    (1, 12, 10, 6, 3, 2, 1, 1, 4),     # pad > md: output larger than input
    (1, 7, 7, 33, 1, 5, 1, 2, 5),      # md not divisible by s2 (r = md // s2), C not a multiple of 32
    (1, 6, 6, 3, 5, 1, 3, 1, 3),       # k = 5, s1 = 3
]


@pytest.mark.parametrize("case", CASES)
def test_correlation_restatements_agree(case):
    N, H, W, C, k, md, s1, s2, pad = case
    rng = np.random.default_rng(0)
    a = rng.standard_normal((N, H, W, C)).astype(np.float32)
    b = rng.standard_normal((N, H, W, C)).astype(np.float32)
    c64 = o.c_correlation(a, b, k, md, s1, s2, pad, f64=True)
    n64 = o.np_correlation(a, b, k, md, s1, s2, pad)
    assert np.abs(c64 - n64).max() < 1e-13
    assert_close(o.c_correlation(a, b, k, md, s1, s2, pad), n64, what="c fp32 vs math")
    g = rng.standard_normal(c64.shape).astype(np.float32)
    ga64, gb64 = o.c_correlation_grad(g, a, b, k, md, s1, s2, pad, f64=True)
    nga, ngb = o.np_correlation_grad(g, a, b, k, md, s1, s2, pad)
    assert np.abs(ga64 - nga).max() < 1e-13 and np.abs(gb64 - ngb).max() < 1e-13
    ga, gb = o.c_correlation_grad(g, a, b, k, md, s1, s2, pad)
    assert_close(ga, nga, what="ga")
    assert_close(gb, ngb, what="gb")


def test_correlation_geometry_and_errors():
    assert o.out_shape(48, 64, 1, 20, 1, 2, 20) == (48, 64, 441)      # flownet_c.py:40
    assert o.out_shape(272, 480, 1, 20, 1, 2, 20) == (272, 480, 441)  # 4K conv3
    assert o.out_shape(10, 12, 3, 3, 2, 2, 3) == (4, 5, 9)
    with pytest.raises(ValueError):
        o.out_shape(48, 64, 2, 20, 1, 2, 20)   # even kernel
    with pytest.raises(ValueError):
        o.out_shape(4, 64, 1, 20, 1, 2, 0)     # neighbourhood does not fit


def test_correlation_known_answers():
    rng = np.random.default_rng(1)
    a = rng.standard_normal((1, 6, 7, 16)).astype(np.float32)
    out = o.c_correlation(a, a, 1, 2, 1, 1, 2, f64=True)
    centre = out[..., 12]                      # zero displacement of a 5x5 grid: mean of squares
    assert np.allclose(centre, (a.astype(np.float64) ** 2).mean(-1))
    # channel order: dy is the OUTER index (correlation_kernel.cu.cc:75-76)
    b = np.zeros_like(a)
    b[0, 3, 4] = 1.0                           # delta at (y=3, x=4)
    out = o.c_correlation(np.ones_like(a), b, 1, 2, 1, 1, 2, f64=True)
    # pixel (2,2) sees the delta at displacement dy=+1, dx=+2 -> d = (1+2)*5 + (2+2) = 19
    assert out[0, 2, 2, 19] == pytest.approx(1.0) and out[0, 2, 2].sum() == pytest.approx(1.0)


def test_correlation_grad_is_autograd_adjoint():
    """Backward0/1 are the exact adjoint of the forward."""
    torch.manual_seed(0)
    k, md, s1, s2, pad = 3, 2, 2, 1, 3
    a = torch.randn(1, 7, 8, 4, dtype=torch.float64, requires_grad=True)
    b = torch.randn(1, 7, 8, 4, dtype=torch.float64, requires_grad=True)
    oh, ow, D = o.out_shape(7, 8, k, md, s1, s2, pad)
    r, gw = md // s2, 2 * (md // s2) + 1
    ext = md + k + r * s2 + 1
    A = torch.nn.functional.pad(a, (0, 0, pad + ext, pad + ext, pad + ext, pad + ext))
    B = torch.nn.functional.pad(b, (0, 0, pad + ext, pad + ext, pad + ext, pad + ext))
    outs = []
    for d in range(D):
        dy, dx = (d // gw - r) * s2, (d % gw - r) * s2
        acc = 0
        for j in range(k):
            for i in range(k):
                ys = ext + md + s1 * torch.arange(oh) + j
                xs = ext + md + s1 * torch.arange(ow) + i
                acc = acc + (A[:, ys][:, :, xs] * B[:, ys + dy][:, :, xs + dx]).sum(-1)
        outs.append(acc / (k * k * 4))
    out = torch.stack(outs, -1)
    g = torch.randn_like(out)
    (out * g).sum().backward()
    ga, gb = o.c_correlation_grad(g.numpy().astype(np.float32), a.detach().numpy().astype(np.float32),
                                  b.detach().numpy().astype(np.float32), k, md, s1, s2, pad, f64=True)
    # inputs were rounded to fp32 for the C call; compare with the same rounding applied to autograd's view
    a32 = a.detach().float().double().requires_grad_()
    b32 = b.detach().float().double().requires_grad_()
    A = torch.nn.functional.pad(a32, (0, 0, pad + ext, pad + ext, pad + ext, pad + ext))
    B = torch.nn.functional.pad(b32, (0, 0, pad + ext, pad + ext, pad + ext, pad + ext))
    outs = []
    for d in range(D):
        dy, dx = (d // gw - r) * s2, (d % gw - r) * s2
        acc = 0
        for j in range(k):
            for i in range(k):
                ys = ext + md + s1 * torch.arange(oh) + j
                xs = ext + md + s1 * torch.arange(ow) + i
                acc = acc + (A[:, ys][:, :, xs] * B[:, ys + dy][:, :, xs + dx]).sum(-1)
        outs.append(acc / (k * k * 4))
    (torch.stack(outs, -1) * g.float().double()).sum().backward()
    assert np.abs(ga - a32.grad.numpy()).max() < 1e-12
    assert np.abs(gb - b32.grad.numpy()).max() < 1e-12


def _warp_case(seed=1, N=2, H=13, W=17, C=3):
    rng = np.random.default_rng(seed)
    img = rng.random((N, H, W, C)).astype(np.float32)
    flow = (rng.standard_normal((N, H, W, 2)) * 4).astype(np.float32)
    flow[0, 0, 0] = [np.nan, 0]
    flow[0, 1, 1] = [100, 0]
    flow[0, 2, 2] = [1.0, -1.0]
    flow[1, 5, 5] = [W - 1 - 5, H - 1 - 5]      # lands exactly on the last pixel: clamped taps
    flow[1, 6, 6] = [-6.5, 0]                   # -1 < x2 < 0: zeroed, not blended
    g = rng.standard_normal((N, H, W, C)).astype(np.float32)
    return img, flow, g


def test_flow_warp_restatements_agree():
    img, flow, g = _warp_case()
    assert_close(o.c_flow_warp(img, flow), o.np_flow_warp(img, flow), what="warp")
    gi, gf = o.c_flow_warp_grad(img, flow, g)
    ngi, ngf = o.np_flow_warp_grad(img, flow, g)
    assert_close(gi, ngi, rel=1e-5, abs_=2e-6, what="gimage")
    assert_close(gf, ngf, rel=1e-5, abs_=2e-6, what="gflow")


def test_flow_warp_known_answers():
    rng = np.random.default_rng(2)
    img = rng.random((1, 6, 8, 3)).astype(np.float32)
    zero = np.zeros((1, 6, 8, 2), np.float32)
    assert np.array_equal(o.c_flow_warp(img, zero), img)                   # zero flow = identity
    shift = zero.copy()
    shift[..., 0] = 2.0                                                    # integer flow = shift, zero fill
    w = o.c_flow_warp(img, shift)
    assert np.array_equal(w[:, :, :6], img[:, :, 2:]) and (w[:, :, 6:] == 0).all()
    nanf = zero.copy()
    nanf[0, 1, 1, 1] = np.nan
    assert (o.c_flow_warp(img, nanf)[0, 1, 1] == 0).all()                  # NaN flow -> 0


def test_flow_warp_flow_grad_is_derivative():
    """the flow gradient is the analytic derivative of the forward (away from knots)."""
    rng = np.random.default_rng(3)
    img = rng.random((1, 9, 9, 2)).astype(np.float32)
    flow = (rng.random((1, 9, 9, 2)) * 0.8 + 0.1).astype(np.float32)       # fractional parts well inside (0,1)
    flow[0, :, 7:, 0] = -0.5
    flow[0, 7:, :, 1] = -0.5                                               # keep all samples in range
    g = rng.standard_normal(img.shape).astype(np.float32)
    _, gf = o.np_flow_warp_grad(img, flow, g)
    eps = 1e-3
    for (y, x, ch) in [(2, 3, 0), (4, 4, 1), (7, 1, 0), (1, 7, 1)]:
        fp, fm = flow.copy(), flow.copy()
        fp[0, y, x, ch] += eps
        fm[0, y, x, ch] -= eps
        num = ((o.np_flow_warp(img, fp) - o.np_flow_warp(img, fm)) * g).sum() / (2 * eps)
        assert num == pytest.approx(gf[0, y, x, ch], rel=1e-3, abs=1e-4)


def test_downsample_restatements_agree_and_known_answers():
    rng = np.random.default_rng(4)
    x = (rng.standard_normal((2, 48, 64, 2)) * 5).astype(np.float32)
    # same size ~ identity, but only approximately: srcx = (dx/(oW-1))*(iW-1) is not exactly dx in float32, so a
    # sliver of weight leaks to the neighbours (downsample_kernel_gpu.cu.cc:40,59)
    assert np.abs(o.c_downsample(x, (48, 64)) - x).max() < 2e-4
    x[0, 10:20, 10:30] = np.nan                                            # NaN block: majority rule
    for size in [(6, 8), (12, 16), (24, 32), (48, 64), (5, 7), (2, 2)]:
        c, n = o.c_downsample(x, size), o.np_downsample(x, size)
        assert_close(c, n, rel=1e-5, abs_=5e-6, what=str(size))
    assert np.isnan(o.c_downsample(x, (24, 32))).any()
    with pytest.raises(ValueError):
        o.c_downsample(x, (1, 8))                                          # reference divides by zero


def test_downsample_rounds_half_away_from_zero():
    # iW=6 -> oW=3: srcx = [0, 2.5, 5] -> isrcx = [0, 3, 5] (np.round would give 2)
    x = np.zeros((1, 2, 6, 1), np.float32)
    x[0, :, 4, 0] = 1.0                       # only reachable from isrcx = 3 (window 3 +- 3, weight 1-1.5/2.5)
    out = o.c_downsample(x, (2, 3))
    assert out[0, 0, 1, 0] > 0
    assert_close(out, o.np_downsample(x, (2, 3)), what="half-away")
