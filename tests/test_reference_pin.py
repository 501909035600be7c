"""Pins the oracle to the REAL reference: the reference's own CUDA kernels (oracle/_ref/libfn2ref.so, compiled
unmodified from /root/reference/src/ops/**.cu.cc by oracle/Makefile) run here on the GPU and are compared with the
oracle and with our kernels on the same inputs.  /root/reference itself is not needed at run time."""
import numpy as np
import pytest
import torch

import flownet2_tf_b200 as fn2
from oracle import oracle as o
from oracle import ref
from conftest import assert_close, corr_inputs, grad_like, warp_inputs

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")]

# Shapes are chosen with H*W % 16 == 0 (or pad >= 16): otherwise the reference's PadData takes its early-return
# branch (pad.cu.cc:31-33) and writes zeros at un-padded coordinates of the padded tensor, racing with / clobbering
# real data when pad < 16 -- a defect of the reference that is deliberately not reproduced.
CASES = [
    (2, 8, 12, 8, 1, 4, 1, 2, 4),
    (1, 8, 12, 5, 3, 3, 2, 2, 3),
    (1, 12, 12, 6, 3, 2, 1, 1, 4),
    (1, 16, 16, 32, 1, 4, 1, 1, 4),
    (1, 14, 18, 32, 1, 20, 1, 2, 20),
    (1, 24, 32, 256, 1, 20, 1, 2, 20),
    (1, 48, 64, 256, 1, 20, 1, 2, 20),
]


@pytest.mark.parametrize("case", CASES)
def test_reference_correlation_vs_oracle_and_ours(case):
    N, H, W, C, k, md, s1, s2, pad = case
    a, b = corr_inputs(N, H, W, C)
    ad, bd = a.cuda(), b.cuda()
    r_out = ref.correlation(ad, bd, k, md, s1, s2, pad)
    torch.cuda.synchronize()
    c_out = o.c_correlation(a.numpy(), b.numpy(), k, md, s1, s2, pad)
    assert_close(c_out, r_out.cpu().numpy(), what="oracle vs reference kernel (forward)")
    assert_close(fn2.correlation(ad, bd, k, md, s1, s2, pad).cpu().numpy(), r_out.cpu().numpy(),
                 what="ours vs reference kernel (forward)")
    g = grad_like(r_out.shape)
    r_ga, r_gb = ref.correlation_grad(g.cuda(), ad, bd, k, md, s1, s2, pad)
    torch.cuda.synchronize()
    c_ga, c_gb = o.c_correlation_grad(g.numpy(), a.numpy(), b.numpy(), k, md, s1, s2, pad)
    assert_close(c_ga, r_ga.cpu().numpy(), what="oracle vs reference (backprops_a)")
    assert_close(c_gb, r_gb.cpu().numpy(), what="oracle vs reference (backprops_b)")
    ga, gb = fn2.correlation_grad(g.cuda(), ad, bd, k, md, s1, s2, pad)
    assert_close(ga.cpu().numpy(), r_ga.cpu().numpy(), what="ours vs reference (backprops_a)")
    assert_close(gb.cpu().numpy(), r_gb.cpu().numpy(), what="ours vs reference (backprops_b)")


@pytest.mark.parametrize("shape", [(2, 13, 17, 3), (1, 40, 70, 3), (1, 9, 33, 2), (1, 8, 8, 40)])
def test_reference_flow_warp_vs_oracle_and_ours(shape):
    N, H, W, C = shape
    image, flow = warp_inputs(N, H, W, C)
    g = grad_like(image.shape)
    im, fl, gd = image.cuda(), flow.cuda(), g.cuda()
    r = ref.flow_warp(im, fl)
    torch.cuda.synchronize()
    assert_close(o.c_flow_warp(image.numpy(), flow.numpy()), r.cpu().numpy(), what="oracle vs reference warp")
    assert_close(fn2.flow_warp(im, fl).cpu().numpy(), r.cpu().numpy(), what="ours vs reference warp")
    r_gi, r_gf = ref.flow_warp_grad(im, fl, gd)
    torch.cuda.synchronize()
    c_gi, c_gf = o.c_flow_warp_grad(image.numpy(), flow.numpy(), g.numpy())
    assert_close(c_gi, r_gi.cpu().numpy(), rel=1e-5, abs_=2e-5, what="oracle vs reference image_grad")
    assert_close(c_gf, r_gf.cpu().numpy(), rel=1e-5, abs_=2e-6, what="oracle vs reference flow_grad")
    gi, gf = fn2.flow_warp_grad(im, fl, gd)
    assert_close(gi.cpu().numpy(), r_gi.cpu().numpy(), rel=1e-5, abs_=2e-5, what="ours vs reference image_grad")
    assert_close(gf.cpu().numpy(), r_gf.cpu().numpy(), rel=1e-5, abs_=2e-6, what="ours vs reference flow_grad")


@pytest.mark.parametrize("size", [(6, 8), (12, 16), (24, 32), (48, 64), (96, 128)])
def test_reference_downsample_vs_oracle_and_ours(size):
    g = torch.Generator().manual_seed(1237)
    x = torch.randn(1, 384, 512, 2, generator=g) * 5
    x[0, 100:140, 200:300] = float("nan")
    r = ref.downsample(x.cuda(), size)
    torch.cuda.synchronize()
    assert_close(o.c_downsample(x.numpy(), size), r.cpu().numpy(), rel=1e-5, abs_=5e-6, what="oracle vs reference")
    assert_close(fn2.downsample(x.cuda(), list(size)).cpu().numpy(), r.cpu().numpy(), rel=1e-5, abs_=5e-6,
                 what="ours vs reference")
