"""GPU parity tests: the CUDA path, called through the Python drop-in API -> ctypes -> C ABI, against the oracle on
identical seeded inputs.  Tolerance for fp32: |err| <= 1e-6 + 1e-5 |ref| (BASELINE.json north_star)."""
import numpy as np
import pytest
import torch

import flownet2_tf_b200 as fn2
from oracle import oracle as o
from conftest import assert_close, corr_inputs, grad_like, warp_inputs

pytestmark = pytest.mark.gpu

CORR_CASES = [
    # N  H   W   C   k  md s1 s2 pad
    (2, 9, 11, 8, 1, 4, 1, 2, 4),
    (1, 10, 12, 5, 3, 3, 2, 2, 3),
    (1, 8, 9, 4, 1, 3, 1, 1, 1),
    (1, 12, 10, 6, 3, 2, 1, 1, 4),
    (1, 7, 7, 33, 1, 5, 1, 2, 5),
    (1, 6, 6, 3, 5, 1, 3, 1, 3),
    (2, 13, 18, 64, 1, 20, 1, 2, 20),   # FlowNetC arguments on a ragged size
    (1, 24, 32, 256, 1, 20, 1, 2, 20),  # FlowNetC arguments, conv3 channel count, quarter size
    (3, 17, 9, 32, 1, 4, 1, 1, 4),      # PWC-style 9x9 window
    (1, 48, 64, 256, 1, 20, 1, 2, 20),  # cfg1: conv3 of one 512x384 pair (flownet_c.py:40)
]


def _corr_case(case, precision="fp32"):
    N, H, W, C, k, md, s1, s2, pad = case
    a, b = corr_inputs(N, H, W, C)
    ad, bd = a.cuda().requires_grad_(), b.cuda().requires_grad_()
    out = fn2.correlation(ad, bd, k, md, s1, s2, pad, precision=precision)
    g = grad_like(out.shape)
    out.backward(g.cuda())
    return a, b, g, out.detach().cpu(), ad.grad.cpu(), bd.grad.cpu()


@pytest.mark.parametrize("case", CORR_CASES)
def test_correlation_fp32_matches_oracle(case):
    N, H, W, C, k, md, s1, s2, pad = case
    a, b, g, out, ga, gb = _corr_case(case)
    want = o.c_correlation(a.numpy(), b.numpy(), k, md, s1, s2, pad, f64=True)
    assert_close(out.numpy(), want, what="forward")
    wga, wgb = o.c_correlation_grad(g.numpy(), a.numpy(), b.numpy(), k, md, s1, s2, pad, f64=True)
    assert_close(ga.numpy(), wga, what="backprops_a")
    assert_close(gb.numpy(), wgb, what="backprops_b")
    # and against the order-faithful fp32 restatement (what the reference kernel itself computes)
    assert_close(out.numpy(), o.c_correlation(a.numpy(), b.numpy(), k, md, s1, s2, pad), what="forward vs fp32 oracle")


def test_correlation_full_size_properties():
    """cfg3 shape (N=32 pairs of 512x384 at conv3): size-independent properties instead of a CPU oracle pass."""
    N, H, W, C = 32, 48, 64, 256
    a, b = corr_inputs(N, H, W, C)
    a, b = a.cuda(), b.cuda()
    args = (1, 20, 1, 2, 20)
    out = fn2.correlation(a, b, *args)
    assert out.shape == (N, 48, 64, 441)
    # (1) batch independence: sample n alone gives the same bits
    assert torch.equal(fn2.correlation(a[5:6].contiguous(), b[5:6].contiguous(), *args)[0], out[5])
    # (2) exact homogeneity under power-of-two scaling
    assert torch.equal(fn2.correlation(a * 2, b * 0.5, *args), out)
    # (3) zero-displacement channel of corr(a, a) is the mean of squares
    aa = fn2.correlation(a, a, *args)
    torch.testing.assert_close(aa[..., 220], (a.double() ** 2).mean(-1).float(), rtol=1e-5, atol=1e-6)
    # (4) swap symmetry: corr(a,b)[y,x,(p,o)] == corr(b,a)[y+2p, x+2o, (-p,-o)]
    ba = fn2.correlation(b, a, *args)
    for (p, q) in [(-10, 10), (3, -4), (0, 7)]:
        d, dm = (p + 10) * 21 + (q + 10), (-p + 10) * 21 + (-q + 10)
        ys = slice(max(0, -2 * p), min(H, H - 2 * p))
        xs = slice(max(0, -2 * q), min(W, W - 2 * q))
        yt = slice(ys.start + 2 * p, ys.stop + 2 * p)
        xt = slice(xs.start + 2 * q, xs.stop + 2 * q)
        torch.testing.assert_close(out[:, ys, xs, d], ba[:, yt, xt, dm], rtol=1e-5, atol=1e-6)
    # (5) gradient = adjoint: <corr(a,b), g> differentiated numerically along a random direction
    g = grad_like(out.shape).cuda()
    ga, gb = fn2.correlation_grad(g, a, b, *args)
    da = torch.randn_like(a)
    lhs = (fn2.correlation(da, b, *args).double() * g.double()).sum()   # the op is linear in a
    rhs = (ga.double() * da.double()).sum()
    assert abs(lhs - rhs) <= 1e-5 * abs(rhs) + 1e-3
    db = torch.randn_like(b)
    lhs = (fn2.correlation(a, db, *args).double() * g.double()).sum()
    rhs = (gb.double() * db.double()).sum()
    assert abs(lhs - rhs) <= 1e-5 * abs(rhs) + 1e-3


def test_correlation_errors_on_device():
    a = torch.zeros(1, 8, 8, 4, device="cuda")
    with pytest.raises(ValueError, match="kernel_size must be odd"):
        fn2.correlation(a, a, 2, 2, 1, 1, 2)
    with pytest.raises(ValueError, match="fit in input height"):
        fn2.correlation(a, a, 1, 20, 1, 2, 0)
    with pytest.raises(ValueError, match="same shape"):
        fn2.correlation(a, torch.zeros(1, 8, 9, 4, device="cuda"), 1, 2, 1, 1, 2)
    with pytest.raises(ValueError, match="contiguous"):
        fn2.correlation(a.permute(0, 2, 1, 3), a.permute(0, 2, 1, 3), 1, 2, 1, 1, 2)


@pytest.mark.parametrize("shape", [(2, 13, 17, 3), (1, 31, 33, 1), (1, 9, 40, 2), (2, 8, 8, 4), (1, 7, 9, 5),
                                   (1, 6, 6, 64), (8, 384, 512, 3),
                                   # W % 4 == 0 and C <= 4: the shared-memory window kernels (flow_warp_tile.cu);
                                   # several tiles, ragged tile edges, windows clipped at every image border
                                   (2, 40, 64, 3), (1, 70, 132, 3), (1, 33, 8, 1), (2, 37, 72, 2), (1, 64, 128, 4),
                                   (1, 100, 200, 3)])
def test_flow_warp_matches_oracle(shape):
    N, H, W, C = shape
    image, flow = warp_inputs(N, H, W, C)
    g = grad_like(image.shape)
    im, fl = image.cuda().requires_grad_(), flow.cuda().requires_grad_()
    out = fn2.flow_warp(im, fl)
    out.backward(g.cuda())
    assert_close(out.detach().cpu().numpy(), o.c_flow_warp(image.numpy(), flow.numpy()), what="warp forward")
    wgi, wgf = o.c_flow_warp_grad(image.numpy(), flow.numpy(), g.numpy())
    # d/dimage sums up to a few hundred atomics per target in the far-flow clusters: order differs run to run
    assert_close(im.grad.cpu().numpy(), wgi, rel=1e-5, abs_=2e-5, what="image_grad")
    assert_close(fl.grad.cpu().numpy(), wgf, rel=1e-5, abs_=2e-6, what="flow_grad")


def test_flow_warp_known_answers_on_device():
    image, _ = warp_inputs(1, 16, 20, 3)
    im = image.cuda()
    zero = torch.zeros(1, 16, 20, 2, device="cuda")
    assert torch.equal(fn2.flow_warp(im, zero), im)
    shift = zero.clone()
    shift[..., 0] = 3.0
    w = fn2.flow_warp(im, shift)
    assert torch.equal(w[:, :, :17], im[:, :, 3:]) and (w[:, :, 17:] == 0).all()
    with pytest.raises(ValueError, match="same batch size, height and width"):
        fn2.flow_warp(im, torch.zeros(1, 16, 21, 2, device="cuda"))


@pytest.mark.parametrize("C", [2, 1, 3, 8])
@pytest.mark.parametrize("size", [(6, 8), (12, 16), (24, 32), (48, 64), (96, 128)])
def test_downsample_matches_oracle(C, size):
    g = torch.Generator().manual_seed(1237)
    x = torch.randn(2, 384, 512, C, generator=g) * 5
    x[0, 100:140, 200:300] = float("nan")          # a NaN block (majority rule) ...
    x.view(-1)[::97] = float("nan")                # ... and ~1 % scattered NaNs
    got = fn2.downsample(x.cuda(), list(size)).cpu().numpy()
    want = o.c_downsample(x.numpy(), size)
    assert_close(got, want, rel=1e-5, abs_=5e-6, what="downsample %s" % (size,))


def test_downsample_small_and_errors():
    g = torch.Generator().manual_seed(7)
    x = torch.randn(3, 9, 14, 2, generator=g)
    for size in [(2, 2), (9, 14), (5, 3), (13, 20)]:   # incl. upsampling, which the op also defines
        assert_close(fn2.downsample(x.cuda(), size).cpu().numpy(), o.c_downsample(x.numpy(), size),
                     rel=1e-5, abs_=5e-6, what=str(size))
    with pytest.raises(ValueError, match=">= 2"):
        fn2.downsample(x.cuda(), [1, 8])
    xr = x.cuda().requires_grad_()
    with pytest.raises(LookupError):
        fn2.downsample(xr, [4, 4]).sum().backward()
