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


def test_large_batch_matches_oracle():
    x = _input(16, 384, 512, 2, 99, nan=False)
    outs = fn2.downsample_multi(x.cuda(), LOSS_SIZES[:2])
    for size, got in zip(LOSS_SIZES[:2], outs):
        assert_close(got.cpu().numpy(), o.c_downsample(x.numpy(), size), rel=1e-5, abs_=5e-6, what="N=16 %s" % (size,))


@pytest.mark.parametrize("shape,sizes", [
    ((3, 48, 64, 2), [(3, 4), (6, 8), (12, 16)]),            # three bands, one warp strip
    ((1, 100, 130, 2), [(7, 9), (13, 33), (50, 40)]),         # ragged: last band of 4 rows, last strip of 2 pixels
    ((2, 37, 53, 1), [(5, 6), (9, 11)]),                      # width not a multiple of 4: single-scale kernels underneath
    ((2, 64, 96, 1), [(4, 6), (8, 12), (16, 24), (32, 48)]),  # one channel: four pixels per thread
    ((1, 50, 1024, 2), [(5, 9), (20, 300)]),                  # 16 warp strips per row
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


# ---- the loss terms in the same pass (f-3, second half): fn2_multiscale_epe_forward ----------------------------
def _epe_reference(x, preds, scale):
    """flownet_c.py:127-160 + utils.py:209-224 on the CPU: oracle downsample of the materialised flow * scale, then
    sqrt(sum_c (pred - down)^2) summed over pixels / N, in float64."""
    xs = (x * np.float32(scale)).astype(np.float32)
    sums, downs = [], []
    for p in preds:
        d = o.c_downsample(xs, (p.shape[1], p.shape[2]))
        downs.append(d)
        diff = p.numpy().astype(np.float64) - d.astype(np.float64)
        sums.append(np.sqrt((diff ** 2).sum(3)).sum() / x.shape[0])
    return np.array(sums), downs


@pytest.mark.parametrize("C,scale", [(2, 0.05), (2, 20.0), (1, 1.0)])
def test_multiscale_epe_matches_the_graph_nodes(C, scale):
    x = _input(3, 384, 512, C, 77, nan=False)
    g = torch.Generator().manual_seed(78)
    preds = [torch.randn(3, h, w, C, generator=g) for h, w in LOSS_SIZES]
    want, wdowns = _epe_reference(x.numpy(), preds, scale)
    pd = [p.cuda().requires_grad_() for p in preds]
    epe, downs = fn2.multiscale_epe(x.cuda(), pd, scale, return_downsampled=True)
    np.testing.assert_allclose(epe.detach().cpu().numpy(), want, rtol=2e-5)
    for d, w in zip(downs, wdowns):   # absolute part scales with the data (samples are 5 * scale in magnitude)
        assert_close(d.cpu().numpy(), w, rel=1e-5, abs_=5e-6 * max(1.0, scale), what="downsampled in the fused loss")
    # gradient w.r.t. the predictions = autograd through the same formula on the kernel's downsampled tensors
    wts = torch.tensor([0.32, 0.08, 0.02, 0.01, 0.005], device="cuda")
    (epe * wts).sum().backward()
    for p, d, w in zip(pd, downs, wts):
        q = p.detach().clone().requires_grad_()
        (torch.sqrt(((q - d) ** 2).sum(3)).sum() / x.shape[0] * w).backward()
        torch.testing.assert_close(p.grad, q.grad, rtol=1e-5, atol=1e-9)
    # the plain op on the scaled tensor gives the same downsampled values bit for bit (scale folded into the load)
    plain = fn2.downsample_multi((x * scale).cuda(), LOSS_SIZES)
    for d, q in zip(downs, plain):
        assert torch.equal(d, q)


def test_multiscale_epe_with_nans_and_reproducible():
    x = _input(2, 96, 128, 2, 5)          # NaN block + scattered NaNs in the ground truth
    g = torch.Generator().manual_seed(6)
    sizes = [(6, 8), (12, 16), (24, 32)]
    preds = [torch.randn(2, h, w, 2, generator=g).cuda() for h, w in sizes]
    epe, downs = fn2.multiscale_epe(x.cuda(), preds, 0.05, return_downsampled=True)
    want, wdowns = _epe_reference(x.numpy(), [p.cpu() for p in preds], 0.05)
    for d, w in zip(downs, wdowns):
        assert_close(d.cpu().numpy(), w, rel=1e-5, abs_=5e-6, what="NaN-majority outputs of the fused loss")
    got = epe.cpu().numpy()
    assert (np.isnan(got) == np.isnan(want)).all()          # a NaN output makes that scale's sum NaN, as in TF
    epe2 = fn2.multiscale_epe(x.cuda(), preds, 0.05)
    assert torch.equal(torch.nan_to_num(epe, nan=-1.0), torch.nan_to_num(epe2, nan=-1.0))


@pytest.mark.parametrize("shape,sizes", [
    ((1, 436, 1024, 2), [(7, 16), (14, 32), (28, 64), (55, 128), (109, 256)]),   # Sintel: band wider than 100 KB
    ((2, 97, 131, 2), [(5, 7), (13, 17), (25, 33)]),                             # odd width: W * C % 4 != 0
    ((1, 30, 40, 3), [(3, 4), (6, 8)]),                                          # C = 3
    ((1, 40, 60, 2), [(30, 45), (5, 7)]),                                        # a reduction below 1.5
])
def test_multiscale_epe_outside_the_fused_kernels(shape, sizes):
    """Shapes fn2_multiscale_epe_forward refuses (the C ABI says so up front) still give the loss terms through the
    Python op: downsample_multi on the scaled flow + the end-point-error nodes, differentiable w.r.t. the predictions."""
    import ctypes
    from flownet2_tf_b200 import _lib
    N, H, W, C = shape
    hw = (ctypes.c_int * (2 * len(sizes)))(*[v for s in sizes for v in s])
    assert _lib.lib().fn2_multiscale_epe_supported(hw, len(sizes), N, H, W, C) == 0
    x = _input(*shape, seed=sum(shape), nan=False)
    g = torch.Generator().manual_seed(4)
    preds = [torch.randn(N, h, w, C, generator=g) for h, w in sizes]
    want, wdowns = _epe_reference(x.numpy(), preds, 0.05)
    pd = [p.cuda().requires_grad_() for p in preds]
    epe, downs = fn2.multiscale_epe(x.cuda(), pd, 0.05, return_downsampled=True)
    np.testing.assert_allclose(epe.detach().cpu().numpy(), want, rtol=2e-5)
    for d, w in zip(downs, wdowns):
        assert_close(d.cpu().numpy(), w, rel=1e-5, abs_=5e-6, what="downsampled, unfused loss %s" % (shape,))
    epe.sum().backward()
    for p, d in zip(pd, downs):
        q = p.detach().clone().requires_grad_()
        (torch.sqrt(((q - d) ** 2).sum(3)).sum() / N).backward()
        torch.testing.assert_close(p.grad, q.grad, rtol=1e-5, atol=1e-9)


def test_results_do_not_depend_on_the_batch_they_are_in():
    """The tiling (16-row bands, one strip per warp) and the summation order are fixed by the image shape alone, so an
    image gives the same bits alone and inside a batch of 32 (order used to depend on N)."""
    x = _input(32, 384, 512, 2, 11, nan=False).cuda()
    full = fn2.downsample_multi(x, LOSS_SIZES)
    one = fn2.downsample_multi(x[13:14].contiguous(), LOSS_SIZES)
    for a, b in zip(full, one):
        assert torch.equal(a[13:14], b)


def test_tie_case_half_of_the_window_nan():
    """Integer scale, exactly half of every window's weight NaN: accum_nan / accum_weight == 1 > 0.5 -> NaN, and with a
    quarter NaN the ratio is 1/3 -> a value; neither sits on the threshold, whatever the rounding of the sums."""
    x = torch.ones(1, 33, 33, 2)
    x[:, :, 1::2] = float("nan")                         # every other column
    out = fn2.downsample_multi(x.cuda(), [(9, 9)])[0].cpu().numpy()
    want = o.c_downsample(x.numpy(), (9, 9))
    assert (np.isnan(out) == np.isnan(want)).all()
    assert_close(out, want, rel=1e-5, abs_=5e-6, what="tie case")


# ---- HFEM 'edges' weighting in the fused pass: fn2_multiscale_epe_edges_forward ------------------------------------
@pytest.mark.parametrize("shape,sizes,lam", [
    ((3, 384, 512, 2), LOSS_SIZES, 2.0),                        # the fused kernels
    ((2, 96, 128, 2), [(6, 8), (12, 16), (24, 32)], 0.5),
    ((1, 436, 1024, 2), [(7, 16), (14, 32), (28, 64)], 2.0),    # wider than the fused kernels take: unfused nodes
])
def test_multiscale_epe_with_edge_weights(shape, sizes, lam):
    """epe[s] = sum(epe_s + lambda * epe_s * downsample(edges, size_s)) / N (utils.py:326-334 on the per-scale edge
    maps of flownet_s_interp.py:177-233) against float64 on the oracle's downsampled tensors; gradient w.r.t. the
    predictions against autograd through the same formula."""
    N, H, W, C = shape
    x = _input(*shape, seed=31, nan=False)
    g = torch.Generator().manual_seed(32)
    edges = torch.rand(N, H, W, 1, generator=g)
    preds = [torch.randn(N, h, w, C, generator=g) for h, w in sizes]
    xs = (x.numpy() * np.float32(0.05)).astype(np.float32)
    want = []
    for p in preds:
        size = (p.shape[1], p.shape[2])
        d = o.c_downsample(xs, size).astype(np.float64)
        e = o.c_downsample(edges.numpy(), size).astype(np.float64)
        epe = np.sqrt(((p.numpy().astype(np.float64) - d) ** 2).sum(3, keepdims=True))
        want.append((epe + lam * epe * e).sum() / N)
    pd = [p.cuda().requires_grad_() for p in preds]
    epe, downs = fn2.multiscale_epe(x.cuda(), pd, 0.05, return_downsampled=True, edges=edges.cuda(), lambda_w=lam)
    np.testing.assert_allclose(epe.detach().cpu().numpy(), want, rtol=2e-5)
    plain = fn2.multiscale_epe(x.cuda(), [p.detach() for p in pd], 0.05)
    assert (epe.detach() > plain).all()                         # the weighting is on
    emaps = fn2.downsample_multi(edges.cuda(), sizes)
    epe.sum().backward()
    for p, d, e in zip(pd, downs, emaps):
        q = p.detach().clone().requires_grad_()
        m = torch.sqrt(((q - d) ** 2).sum(3, keepdim=True))
        ((m + lam * (m * e)).sum() / N).backward()
        torch.testing.assert_close(p.grad, q.grad, rtol=1e-5, atol=1e-9)
    with pytest.raises(ValueError):
        fn2.multiscale_epe(x.cuda(), [p.detach() for p in pd], 0.05, edges=edges[:, :-1].cuda())
