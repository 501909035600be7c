"""fn2_correlation_forward_fused: LeakyReLU + concat-buffer store in the correlation's
epilogue must equal the three separate graph nodes of flownet_c.py:40-46, forward and backward."""
import numpy as np
import pytest
import torch

from conftest import corr_inputs, assert_close

pytestmark = pytest.mark.gpu

import flownet2_tf_b200 as fn2  # noqa: E402
from oracle import oracle as o  # noqa: E402

ARGS = (1, 20, 1, 2, 20)


@pytest.mark.parametrize("precision", ["tf32", "fp32"])
@pytest.mark.parametrize("shape", [(2, 48, 64, 256), (1, 13, 21, 64), (1, 24, 160, 32)])
def test_fused_equals_unfused_nodes(precision, shape):
    a, b = corr_inputs(*shape)
    a, b = (a - 0.3).cuda(), b.cuda()                       # both signs in the volume
    N, H, W, _ = shape
    redir = torch.randn(N, H, W, 32, device="cuda")
    buf = torch.full((N, H, W, 32 + 441 + 7), 123.0, device="cuda")
    buf[..., :32] = redir
    out = fn2.correlation_leaky_concat(a, b, *ARGS, buf, 32, 0.1, precision=precision)
    assert out.data_ptr() == buf.data_ptr()
    cc = fn2.correlation(a, b, *ARGS, precision=precision)
    want = torch.nn.functional.leaky_relu(cc, 0.1)
    assert torch.equal(out[..., 32:473], want)               # same arithmetic, same bits
    assert torch.equal(out[..., :32], redir)                 # neighbours untouched
    assert bool((out[..., 473:] == 123.0).all())
    assert bool((want < 0).any()) and bool((want > 0).any())


def test_fused_matches_oracle_fp32():
    a, b = corr_inputs(1, 12, 16, 64)
    a = a - 0.3
    buf = torch.zeros(1, 12, 16, 473, device="cuda")
    fn2.correlation_leaky_concat(a.cuda(), b.cuda(), *ARGS, buf, 32, 0.1, precision="fp32")
    cc = o.c_correlation(a.numpy(), b.numpy(), *ARGS, f64=True)
    want = np.where(cc > 0, cc, 0.1 * cc)
    assert_close(buf[..., 32:].cpu().numpy(), want, rel=1e-5, abs_=1e-6, what="fused fp32")


@pytest.mark.parametrize("precision,shape", [("tf32", (2, 24, 32, 128)), ("tf32x3", (2, 24, 32, 128)),
                                             ("tf32x3", (1, 13, 70, 64)), ("fp32", (1, 10, 12, 32))])
def test_fused_backward_equals_unfused_backward(precision, shape):
    """fn2_correlation_backward_fused (concat slice + LeakyReLU derivative inside the CorrelationGrad kernel's gradient
    gather) against the three separate graph nodes: the same products in the same order, so bit-equal."""
    N, H, W, C = shape
    a0, b0 = corr_inputs(N, H, W, C)
    a0 = a0 - 0.3
    g = torch.randn(N, H, W, 473, generator=torch.Generator().manual_seed(3)).cuda()
    res = []
    for fused in (True, False):
        a, b = a0.cuda().requires_grad_(), b0.cuda().requires_grad_()
        redir = torch.randn(N, H, W, 32, generator=torch.Generator().manual_seed(4)).cuda().requires_grad_()
        if fused:
            buf = torch.empty(N, H, W, 473, device="cuda")
            buf[..., :32] = redir
            net = fn2.correlation_leaky_concat(a, b, *ARGS, buf, 32, 0.1, precision=precision)
        else:
            cc = torch.nn.functional.leaky_relu(fn2.correlation(a, b, *ARGS, precision=precision), 0.1)
            net = torch.cat([redir, cc], dim=3)
        (net * g).sum().backward()
        res.append((a.grad.clone(), b.grad.clone(), redir.grad.clone()))
    for x, y in zip(*res):
        assert torch.equal(x, y)


def test_fused_backward_matches_oracle():
    """The fused backward against the float64 oracle of CorrelationGrad fed with the masked, sliced gradient."""
    a, b = corr_inputs(1, 12, 16, 64)
    a = a - 0.2
    buf = torch.zeros(1, 12, 16, 473, device="cuda")
    ad, bd = a.cuda().requires_grad_(), b.cuda().requires_grad_()
    net = fn2.correlation_leaky_concat(ad, bd, *ARGS, buf, 32, 0.1, precision="tf32x3")
    g = torch.randn(1, 12, 16, 473, generator=torch.Generator().manual_seed(5))
    (net * g.cuda()).sum().backward()
    cc = o.c_correlation(a.numpy(), b.numpy(), *ARGS, f64=True)
    gm = (g[..., 32:].numpy() * np.where(cc > 0, 1.0, 0.1)).astype(np.float32)
    wga, wgb = o.c_correlation_grad(gm, a.numpy(), b.numpy(), *ARGS, f64=True)
    assert_close(ad.grad.cpu().numpy(), wga, what="fused backprops_a")
    assert_close(bd.grad.cpu().numpy(), wgb, what="fused backprops_b")


def test_fused_errors():
    a, b = corr_inputs(1, 8, 8, 32)
    a, b = a.cuda(), b.cuda()
    with pytest.raises(ValueError, match="cannot hold"):
        fn2.correlation_leaky_concat(a, b, *ARGS, torch.empty(1, 8, 8, 440, device="cuda"), 0)
    with pytest.raises(ValueError, match="multiple of 4"):
        fn2.correlation_leaky_concat(a, b, *ARGS, torch.empty(1, 8, 8, 473, device="cuda"), 30)
    with pytest.raises(ValueError):
        fn2.correlation_leaky_concat(a, b, *ARGS, torch.empty(1, 8, 8, 473), 32)


@pytest.mark.parametrize("shape", [(2, 384, 512, 3), (1, 37, 52, 3), (1, 20, 30, 3), (1, 16, 24, 8), (2, 33, 64, 1)])
def test_flow_warp_error_equals_warp_then_norm(shape):
    from conftest import warp_inputs
    N, H, W, C = shape
    image, flow = warp_inputs(N, H, W, C)
    ref = torch.rand(N, H, W, C, generator=torch.Generator().manual_seed(11))
    image, flow, ref = image.cuda(), flow.cuda(), ref.cuda()
    warped, err = fn2.flow_warp_error(image, flow, ref)
    want_w = fn2.flow_warp(image, flow)
    assert torch.equal(warped, want_w)
    want_e = torch.sqrt(((ref - want_w) ** 2).sum(3, keepdim=True))
    torch.testing.assert_close(err, want_e, rtol=2e-6, atol=1e-7)
    assert_close(warped.cpu().numpy(), o.c_flow_warp(image.cpu().numpy(), flow.cpu().numpy()), what="warp")
    with pytest.raises(LookupError):
        fn2.flow_warp_error(image.clone().requires_grad_(), flow, ref)
