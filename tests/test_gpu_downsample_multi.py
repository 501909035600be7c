"""fn2_downsample_multi_forward: every scale of the fused pass against the CPU oracle of the
single-size op (downsample_kernel_gpu.cu.cc:35-76) and against the single-size kernel."""
import numpy as np
import pytest
import torch

from conftest import assert_close

pytestmark = pytest.mark.gpu

import flownet2_tf_b200 as fn2  # noqa: E402
from oracle import oracle as o  # noqa: E402

LOSS_SIZES = [(6, 8), (12, 16), (24, 32), (48, 64), (96, 128)]   # flownet_c.py:135-160


def _input(N, H, W, C, seed, nan=True):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(N, H, W, C, generator=g) * 5
    if nan:
        x[0, H // 4:H // 4 + H // 9, W // 3:W // 3 + W // 5] = float("nan")   # a NaN block (majority rule) ...
        x.view(-1)[::97] = float("nan")                                      # ... and ~1 % scattered NaNs
    return x


@pytest.mark.parametrize("C", [2, 1])
def test_loss_scales_match_oracle(C):
    x = _input(2, 384, 512, C, 1237)
    outs = fn2.downsample_multi(x.cuda(), LOSS_SIZES)
    assert [tuple(t.shape) for t in outs] == [(2, h, w, C) for h, w in LOSS_SIZES]
    for size, got in zip(LOSS_SIZES, outs):
        want = o.c_downsample(x.numpy(), size)
        assert_close(got.cpu().numpy(), want, rel=1e-5, abs_=5e-6, what="multi %s C=%d" % (size, C))


@pytest.mark.parametrize("ranks", [8, 16])
def test_both_cluster_sizes_match_oracle(ranks, monkeypatch):
    """Small batches take 16-CTA clusters, batches that fill the chip 8 (downsample_multi.cu); the developer switch
    forces either, so both row partitions are checked at the loss shape whatever the batch."""
    monkeypatch.setenv("FN2_DS_MAX_RANKS", str(ranks))
    x = _input(2, 384, 512, 2, 4321)
    outs = fn2.downsample_multi(x.cuda(), LOSS_SIZES)
    for size, got in zip(LOSS_SIZES, outs):
        assert_close(got.cpu().numpy(), o.c_downsample(x.numpy(), size), rel=1e-5, abs_=5e-6,
                     what="ranks %d %s" % (ranks, size))


def test_large_batch_takes_the_portable_cluster_size():
    """More than ~1.3 CTAs per SM with 16 ranks: the launch takes 8 ranks per image; same numbers as the oracle."""
    x = _input(16, 384, 512, 2, 99, nan=False)
    outs = fn2.downsample_multi(x.cuda(), LOSS_SIZES[:2])
    for size, got in zip(LOSS_SIZES[:2], outs):
        assert_close(got.cpu().numpy(), o.c_downsample(x.numpy(), size), rel=1e-5, abs_=5e-6, what="N=16 %s" % (size,))


@pytest.mark.parametrize("shape,sizes", [
    ((3, 48, 64, 2), [(3, 4), (6, 8), (12, 16)]),            # two cluster ranks
    ((1, 100, 130, 2), [(7, 9), (13, 33), (50, 40)]),         # ragged: rows not divisible by the ranks, odd widths
    ((2, 37, 53, 1), [(5, 6), (9, 11)]),                      # one pixel per thread, single rank
    ((2, 64, 96, 1), [(4, 6), (8, 12), (16, 24), (32, 48)]),  # four floats per thread
    ((1, 384, 512, 2), [(6, 8)]),                             # a single scale through the fused kernel
    ((1, 40, 60, 2), [(2, 2), (3, 5), (5, 7), (9, 9), (11, 13), (13, 17), (20, 30)]),   # more than five scales
    ((2, 30, 40, 3), [(3, 4), (6, 8)]),                       # C = 3: single-scale kernels underneath
    ((1, 9, 14, 2), [(9, 14), (13, 20), (2, 2)]),             # identity / upsampling: single-scale kernels underneath
])
def test_shapes_match_oracle_and_single_scale_op(shape, sizes):
    x = _input(*shape, seed=sum(shape))
    xd = x.cuda()
    outs = fn2.downsample_multi(xd, sizes)
    for size, got in zip(sizes, outs):
        assert_close(got.cpu().numpy(), o.c_downsample(x.numpy(), size), rel=1e-5, abs_=5e-6, what=str((shape, size)))
        assert_close(got.cpu().numpy(), fn2.downsample(xd, list(size)).cpu().numpy(), rel=2e-6, abs_=2e-6,
                     what="vs single %s" % (size,))


def test_all_nan_and_constant():
    x = torch.full((1, 96, 128, 2), 3.25)
    outs = fn2.downsample_multi(x.cuda(), [(6, 8), (24, 32)])
    for t in outs:   # a weighted average of a constant is the constant
        np.testing.assert_allclose(t.cpu().numpy(), 3.25, rtol=2e-6)
    x[:] = float("nan")
    for t in fn2.downsample_multi(x.cuda(), [(6, 8), (24, 32)]):
        assert np.isnan(t.cpu().numpy()).all()


def test_deterministic_and_graph_capturable():
    x = _input(4, 384, 512, 2, 5).cuda()
    a = [t.clone() for t in fn2.downsample_multi(x, LOSS_SIZES)]
    b = fn2.downsample_multi(x, LOSS_SIZES)
    for s, t in zip(a, b):
        assert torch.equal(torch.nan_to_num(s, nan=-7.0), torch.nan_to_num(t, nan=-7.0))
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        fn2.downsample_multi(x, LOSS_SIZES)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            c = fn2.downsample_multi(x, LOSS_SIZES)
        gr.replay()
    torch.cuda.synchronize()
    for s, t in zip(a, c):
        assert torch.equal(torch.nan_to_num(s, nan=-7.0), torch.nan_to_num(t, nan=-7.0))


def test_errors():
    x = torch.randn(1, 20, 30, 2).cuda()
    with pytest.raises(ValueError, match=">= 2"):
        fn2.downsample_multi(x, [(4, 4), (1, 8)])
    with pytest.raises(ValueError, match="2 dimensions"):
        fn2.downsample_multi(x, [(4, 4, 4)])
    with pytest.raises(ValueError):
        fn2.downsample_multi(x.cpu(), [(4, 4)])
    with pytest.raises(LookupError):
        fn2.downsample_multi(x.clone().requires_grad_(), [(4, 4)])
    assert fn2.downsample_multi(x, []) == []
