"""GPU parity tests of the tensor-core (tcgen05) correlation paths against the f64 oracle.

Stated tolerances (BASELINE.json north_star: "the tf32/bf16 paths within a stated tolerance"):
  tf32   : |err| <= 2.5e-3 * max|ref|   (10-bit mantissa operands, fp32 accumulate over C = 256 products; measured
           worst case on these inputs is ~9e-4 * max|ref|, profiles/r01_b_check_tc_bwd.log)
  tf32x3 : the fp32 tolerance of the op, 1e-5 relative / 1e-6 absolute (split-tf32, fp32-grade): checked with
           conftest.assert_close exactly like the fp32 CUDA-core kernels
The same figure is stated in BASELINE.md section 4 and DESIGN.md section 2.
"""
import numpy as np
import pytest
import torch

import flownet2_tf_b200 as fn2
from oracle import oracle as o
from conftest import assert_close, corr_inputs, grad_like

pytestmark = pytest.mark.gpu

ARGS = (1, 20, 1, 2, 20)
TF32_TOL = 2.5e-3

TC_SHAPES = [
    (1, 48, 64, 256),   # cfg1: conv3 of one 512x384 pair
    (2, 13, 18, 64),    # ragged: odd class sizes, one channel half
    (3, 47, 61, 96),    # ragged, three 32-channel chunks
    (1, 24, 32, 256),
    (2, 5, 7, 32),      # smaller than one tile
    (1, 1, 64, 32),     # a single row: the y%2 == 1 classes are empty
]


def _rel_to_max(got, want):
    return float(np.abs(got.astype(np.float64) - want).max() / max(np.abs(want).max(), 1e-30))


@pytest.mark.parametrize("shape", TC_SHAPES)
def test_tf32_forward_backward_matches_oracle(shape):
    N, H, W, C = shape
    a, b = corr_inputs(N, H, W, C)
    ad, bd = a.cuda().requires_grad_(), b.cuda().requires_grad_()
    out = fn2.correlation(ad, bd, *ARGS, precision="tf32")
    g = grad_like(out.shape)
    out.backward(g.cuda())
    want = o.c_correlation(a.numpy(), b.numpy(), *ARGS, f64=True)
    assert _rel_to_max(out.detach().cpu().numpy(), want) <= TF32_TOL
    wga, wgb = o.c_correlation_grad(g.numpy(), a.numpy(), b.numpy(), *ARGS, f64=True)
    assert _rel_to_max(ad.grad.cpu().numpy(), wga) <= TF32_TOL
    assert _rel_to_max(bd.grad.cpu().numpy(), wgb) <= TF32_TOL


@pytest.mark.parametrize("shape", TC_SHAPES + [(2, 20, 140, 64), (1, 8, 70, 160)])   # + wide classes (TX > 1)
def test_tf32x3_forward_backward_meets_the_fp32_tolerance(shape):
    """The split-tf32 tensor-core path is held to the tolerance of the reference's own precision."""
    N, H, W, C = shape
    a, b = corr_inputs(N, H, W, C)
    ad, bd = a.cuda().requires_grad_(), b.cuda().requires_grad_()
    out = fn2.correlation(ad, bd, *ARGS, precision="tf32x3")
    g = grad_like(out.shape)
    out.backward(g.cuda())
    want = o.c_correlation(a.numpy(), b.numpy(), *ARGS, f64=True)
    assert_close(out.detach().cpu().numpy(), want, rel=1e-5, abs_=1e-6, what="tf32x3 forward")
    wga, wgb = o.c_correlation_grad(g.numpy(), a.numpy(), b.numpy(), *ARGS, f64=True)
    assert_close(ad.grad.cpu().numpy(), wga, rel=1e-5, abs_=1e-6, what="tf32x3 backprops_a")
    assert_close(bd.grad.cpu().numpy(), wgb, rel=1e-5, abs_=1e-6, what="tf32x3 backprops_b")


def test_tf32x3_large_dynamic_range_and_specials():
    """Magnitudes from 1e-12 to 1e+12 (hi/lo split must not lose small operands or overflow), and an inf / NaN input
    must come out as inf / NaN in exactly the outputs it touches, not as NaN everywhere (lo of inf is 0, not inf-inf)."""
    N, H, W, C = 1, 12, 40, 64
    g = torch.Generator().manual_seed(11)
    scale = 10.0 ** torch.randint(-6, 7, (N, H, W, 1), generator=g).float()
    a = torch.randn(N, H, W, C, generator=g) * scale
    b = torch.randn(N, H, W, C, generator=g) * scale.flip(2)
    out = fn2.correlation(a.cuda(), b.cuda(), *ARGS, precision="tf32x3").cpu().numpy()
    want = o.c_correlation(a.numpy(), b.numpy(), *ARGS, f64=True)
    # per output the error scales with sum |a_c b_c| / C, which here can dwarf the (cancelling) sum itself
    mag = o.c_correlation(a.abs().numpy(), b.abs().numpy(), *ARGS, f64=True)
    assert (np.abs(out - want) <= 4e-6 * mag + 1e-30).all()
    a2, b2 = corr_inputs(N, H, W, C)
    a2[0, 4, 6, 3] = float("inf")
    b2[0, 6, 20, 5] = float("nan")
    out = fn2.correlation(a2.cuda(), b2.cuda(), *ARGS, precision="tf32x3").cpu().numpy()
    want = o.c_correlation(a2.numpy(), b2.numpy(), *ARGS, f64=True)
    assert (np.isnan(out) == np.isnan(want)).all() and (np.isinf(out) == np.isinf(want)).all()
    ok = np.isfinite(want)
    assert_close(out[ok], want[ok], rel=1e-5, abs_=1e-6, what="tf32x3 finite outputs next to inf/NaN inputs")


@pytest.mark.parametrize("precision,check", [("tf32", "max"), ("tf32x3", "fp32")])
def test_value_check_on_a_full_batch(precision, check):
    """cfg3 batch (N = 32, the benchmarked launch): two pairs picked out of the batch are checked VALUE by value
    against the f64 oracle, forward and both gradients (the other tests at this size check properties only)."""
    N, H, W, C = 32, 48, 64, 256
    a, b = corr_inputs(N, H, W, C)
    g = grad_like((N, H, W, 441))
    ad, bd, gd = a.cuda(), b.cuda(), g.cuda()
    out = fn2.correlation(ad, bd, *ARGS, precision=precision)
    ga, gb = fn2.correlation_grad(gd, ad, bd, *ARGS, precision=precision)
    for n in (5, 29):
        sl = slice(n, n + 1)
        want = o.c_correlation(a[sl].numpy(), b[sl].numpy(), *ARGS, f64=True)
        wga, wgb = o.c_correlation_grad(g[sl].numpy(), a[sl].numpy(), b[sl].numpy(), *ARGS, f64=True)
        for got, ref_, what in ((out[sl], want, "forward"), (ga[sl], wga, "backprops_a"), (gb[sl], wgb, "backprops_b")):
            got = got.cpu().numpy()
            if check == "max":
                assert _rel_to_max(got, ref_) <= TF32_TOL, (what, n)
            else:
                assert_close(got, ref_, rel=1e-5, abs_=1e-6, what="%s %s pair %d" % (precision, what, n))


def test_tf32_exact_on_tf32_representable_inputs():
    """With inputs that are exactly representable in tf32 (and small integers, so every product and partial sum is
    exact in fp32) the tensor-core path must agree with the oracle to the last bit: isolates indexing from rounding."""
    N, H, W, C = 2, 48, 64, 256
    g = torch.Generator().manual_seed(5)
    a = torch.randint(-4, 5, (N, H, W, C), generator=g).float()
    b = torch.randint(-4, 5, (N, H, W, C), generator=g).float()
    out = fn2.correlation(a.cuda(), b.cuda(), *ARGS, precision="tf32")
    want = o.c_correlation(a.numpy(), b.numpy(), *ARGS, f64=True)
    np.testing.assert_array_equal(out.cpu().numpy().astype(np.float64), want)
    go = torch.randint(-4, 5, out.shape, generator=g).float()
    ga, gb = fn2.correlation_grad(go.cuda(), a.cuda(), b.cuda(), *ARGS, precision="tf32")
    wga, wgb = o.c_correlation_grad(go.numpy(), a.numpy(), b.numpy(), *ARGS, f64=True)
    np.testing.assert_array_equal(ga.cpu().numpy().astype(np.float64), wga)
    np.testing.assert_array_equal(gb.cpu().numpy().astype(np.float64), wgb)


def test_tf32_full_size_properties():
    """cfg3 shape (N = 32): batch independence, adjointness of backward and forward, zero-padding rows."""
    N, H, W, C = 32, 48, 64, 256
    a, b = corr_inputs(N, H, W, C)
    a, b = a.cuda(), b.cuda()
    out = fn2.correlation(a, b, *ARGS, precision="tf32")
    assert torch.equal(fn2.correlation(a[7:8].contiguous(), b[7:8].contiguous(), *ARGS, precision="tf32")[0], out[7])
    # displacements that point outside the image are exact zeros (the reference's zero padding)
    assert (out[:, 0, :, :10 * 21] == 0).all() and (out[:, :, 0].reshape(N, H, 21, 21)[..., :10] == 0).all()
    g = grad_like(out.shape).cuda()
    ga, gb = fn2.correlation_grad(g, a, b, *ARGS, precision="tf32")
    ga7, gb7 = fn2.correlation_grad(g[7:8].contiguous(), a[7:8].contiguous(), b[7:8].contiguous(), *ARGS,
                                    precision="tf32")
    assert torch.equal(ga7[0], ga[7]) and torch.equal(gb7[0], gb[7])
    gen = torch.Generator(device="cuda").manual_seed(99)
    da = torch.randn(a.shape, device="cuda", generator=gen)
    db = torch.randn(b.shape, device="cuda", generator=gen)
    # <corr(da, b), g> == <ga, da> and <corr(a, db), g> == <gb, db>.  Both sides round every product to tf32 (and the
    # tensor core truncates), and the totals are cancellation-heavy sums, so the bound is relative to the sum of
    # magnitudes: 2^-10 per operand pair would allow 2e-3; 5e-4 is what independent rounding leaves.
    lhs = (fn2.correlation(da, b, *ARGS, precision="tf32").double() * g.double()).sum()
    prod = ga.double() * da.double()
    assert abs(lhs - prod.sum()) <= 5e-4 * prod.abs().sum()
    lhs = (fn2.correlation(a, db, *ARGS, precision="tf32").double() * g.double()).sum()
    prod = gb.double() * db.double()
    assert abs(lhs - prod.sum()) <= 5e-4 * prod.abs().sum()


def test_unsupported_geometry_falls_back_to_fp32_with_a_warning_never_to_less_accuracy():
    """The C ABI answers FN2_E_UNSUPPORTED_PRECISION (no silent switch inside the library); the Python op then runs the
    fp32 CUDA-core kernels -- MORE accurate than what was asked for -- and says so.  There is still no CPU path."""
    from flownet2_tf_b200 import _lib
    a, b = corr_inputs(1, 8, 8, 4)     # C % 32 != 0: no tensor-core tiling
    ad, bd = a.cuda(), b.cuda()
    out = torch.empty(1, 8, 8, 9, device="cuda")
    for prec in (1, 3):
        assert _lib.lib().fn2_correlation_precision_supported(1, 8, 8, 4, 1, 2, 1, 2, 2, prec) == 0
        rc = _lib.lib().fn2_correlation_forward(ad.data_ptr(), bd.data_ptr(), out.data_ptr(), 1, 8, 8, 4, 1, 2, 1, 2, 2,
                                                prec, None, 0, None)
        assert rc == -7
    assert _lib.lib().fn2_correlation_precision_supported(1, 48, 64, 256, 1, 20, 1, 2, 20, 3) == 1
    with pytest.warns(RuntimeWarning, match="fp32 CUDA-core"):
        got = fn2.correlation(ad, bd, 1, 2, 1, 2, 2, precision="tf32")
    assert torch.equal(got, fn2.correlation(ad, bd, 1, 2, 1, 2, 2, precision="fp32"))
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("error")                      # "auto" picks fp32 here silently
        assert torch.equal(fn2.correlation(ad, bd, 1, 2, 1, 2, 2), got)


def test_two_devices_in_one_process():
    """Function attributes (opt-in shared memory, cluster sizes) are per device: a process that first runs on cuda:0
    and then on cuda:1 must set them again there.  Skipped on a one-GPU box."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one process")
    a, b = corr_inputs(2, 24, 32, 64)
    res = []
    for dev in ("cuda:0", "cuda:1"):
        ad, bd = a.to(dev), b.to(dev)
        out = fn2.correlation(ad, bd, *ARGS, precision="tf32x3")
        ga, gb = fn2.correlation_grad(torch.ones_like(out), ad, bd, *ARGS, precision="tf32")
        flow = torch.zeros(2, 24, 32, 2, device=dev)
        w = fn2.flow_warp(ad[..., :3].contiguous(), flow)
        ds = fn2.downsample_multi(torch.randn(2, 64, 64, 2, generator=torch.Generator().manual_seed(1)).to(dev),
                                  [(4, 4), (8, 8)])
        torch.cuda.synchronize(dev)
        res.append([t.cpu() for t in (out, ga, gb, w, *ds)])
    for x, y in zip(*res):
        assert torch.equal(x, y)


@pytest.mark.parametrize("shape", [
    (40, 47, 61, 64),    # more pair tiles than resident pairs, classes of different sizes (invalid tiles in the list),
                         # partial column tiles, a tail that is split
    (75, 24, 32, 32),    # 300 pair tiles on 74 pairs: 4 rounds + a split tail of 4 tiles
    (37, 16, 32, 128),   # exactly two rounds, no tail (C a power of two: the 1/C scaling stays exact)
    (3, 48, 64, 256),    # fewer pair tiles than resident pairs (one tile each)
    (2, 20, 140, 32),    # wide classes: three column tiles, two 32-column b blocks per group of b-rows
])
@pytest.mark.parametrize("precision", ["tf32", "tf32x3"])
def test_persistent_pair_forward_is_exact_on_integer_inputs(shape, precision):
    """The persistent CTA-pair forward walks a work list with running barrier phases; integer inputs make every sum
    exact, so any tile taken twice, skipped, or computed against a stale a tile shows as a wrong integer."""
    N, H, W, C = shape
    g = torch.Generator().manual_seed(sum(shape))
    a = torch.randint(-3, 4, shape, generator=g).float()
    b = torch.randint(-3, 4, shape, generator=g).float()
    out = fn2.correlation(a.cuda(), b.cuda(), *ARGS, precision=precision)
    want = o.c_correlation(a.numpy(), b.numpy(), *ARGS, f64=True)
    np.testing.assert_array_equal(out.cpu().numpy().astype(np.float64), want)
    out2 = fn2.correlation(a.cuda(), b.cuda(), *ARGS, precision=precision)
    assert torch.equal(out, out2)


@pytest.mark.parametrize("shape", [
    (14, 47, 61, 64),    # 336 tiles on 148 CTAs: every CTA walks over several tiles, ragged classes, partial tiles
    (37, 16, 32, 128),   # exactly two tiles per CTA, one row tile per class
    (6, 20, 140, 32),    # wide classes: three column tiles, two 32-pixel items per source row
    (3, 48, 64, 256),    # fewer tiles than CTAs (one tile each)
])
@pytest.mark.parametrize("precision", ["tf32", "tf32x3"])
def test_persistent_backward_is_exact_on_integer_inputs(shape, precision):
    """The persistent backward kernel carries its stage ring, row barriers and accumulator from tile to tile; integer
    inputs make every sum exact (C is a power of two), so a stale G block, an undrained accumulator row or a tile
    taken twice shows as a wrong number."""
    N, H, W, C = shape
    g = torch.Generator().manual_seed(sum(shape))
    a = torch.randint(-3, 4, shape, generator=g).float()
    b = torch.randint(-3, 4, shape, generator=g).float()
    gr = torch.randint(-3, 4, (N, H, W, 441), generator=g).float()
    ga, gb = fn2.correlation_grad(gr.cuda(), a.cuda(), b.cuda(), *ARGS, precision=precision)
    wa, wb = o.c_correlation_grad(gr.numpy(), a.numpy(), b.numpy(), *ARGS, f64=True)
    np.testing.assert_array_equal(ga.cpu().numpy().astype(np.float64), wa)
    np.testing.assert_array_equal(gb.cpu().numpy().astype(np.float64), wb)
    ga2, gb2 = fn2.correlation_grad(gr.cuda(), a.cuda(), b.cuda(), *ARGS, precision=precision)
    assert torch.equal(ga, ga2) and torch.equal(gb, gb2)


def test_first_cuda_action_of_a_fresh_thread_is_one_of_our_launches():
    """PyTorch's autograd worker may reach our backward as its first CUDA action (smoke() does): the tensor-map
    encode is a driver call and needs a context bound to that thread (CUDA_ERROR_INVALID_CONTEXT otherwise)."""
    import threading
    a, b = corr_inputs(1, 12, 16, 64)
    ad, bd = a.cuda(), b.cuda()
    go = grad_like((1, 12, 16, 441)).cuda()
    img, flo = torch.rand(1, 32, 64, 3).cuda(), (torch.randn(1, 32, 64, 2) * 3).cuda()
    torch.cuda.synchronize()
    res = {}

    def work():
        try:
            res["grads"] = fn2.correlation_grad(go, ad, bd, 1, 20, 1, 2, 20, precision="tf32x3")
            res["warp"] = fn2.flow_warp(img, flo)
            torch.cuda.synchronize()
        except Exception as exc:  # pragma: no cover
            res["error"] = exc

    t = threading.Thread(target=work)
    t.start()
    t.join()
    assert "error" not in res, res.get("error")
    wga, wgb = o.c_correlation_grad(go.cpu().numpy(), a.numpy(), b.numpy(), 1, 20, 1, 2, 20, f64=True)
    assert_close(res["grads"][0].cpu().numpy(), wga)
    assert_close(res["grads"][1].cpu().numpy(), wgb)


def test_more_than_2_to_the_31_elements():
    """Maximum sizes: a cost volume of 1600 x 48 x 64 x 441 = 2.17e9 elements (8.7 GB) -- every index past 2^31 must
    be formed in 64 bits.  Images are independent, so a batch made of four distinct pairs repeated 400 times must
    reproduce the four-pair result bit for bit, in the forward and in both gradients, at the start, around the 2^31st
    element (image 1585) and at the end."""
    if torch.cuda.get_device_properties(0).total_memory < 80e9:
        pytest.skip("needs ~45 GB of device memory")
    reps, base_n = 400, 4
    a4, b4 = corr_inputs(base_n, 48, 64, 256)
    g4 = grad_like((base_n, 48, 64, 441))
    a4, b4, g4 = a4.cuda(), b4.cuda(), g4.cuda()
    out4 = fn2.correlation(a4, b4, 1, 20, 1, 2, 20)
    ga4, gb4 = fn2.correlation_grad(g4, a4, b4, 1, 20, 1, 2, 20)
    a, b = a4.repeat(reps, 1, 1, 1), b4.repeat(reps, 1, 1, 1)
    out = fn2.correlation(a, b, 1, 20, 1, 2, 20)
    assert out.numel() > 2 ** 31
    picks = [0, 1, 2, 3, 1584, 1585, 1586, 1587, 1596, 1597, 1598, 1599]
    for n in picks:
        assert torch.equal(out[n], out4[n % base_n]), n
    g = g4.repeat(reps, 1, 1, 1)
    del out
    ga, gb = fn2.correlation_grad(g, a, b, 1, 20, 1, 2, 20)
    for n in picks:
        assert torch.equal(ga[n], ga4[n % base_n]) and torch.equal(gb[n], gb4[n % base_n]), n
    del a, b, g, ga, gb
    # flow_warp: 3700 images of 384 x 512 x 3 = 2.18e9 elements
    img2, flo2 = torch.rand(2, 384, 512, 3, device="cuda"), torch.randn(2, 384, 512, 2, device="cuda") * 8
    w2 = fn2.flow_warp(img2, flo2)
    n_img = 3700
    w = fn2.flow_warp(img2.repeat(n_img // 2, 1, 1, 1), flo2.repeat(n_img // 2, 1, 1, 1))
    assert w.numel() > 2 ** 31
    for n in (0, 1, 3640, 3641, 3642, 3698, 3699):
        assert torch.equal(w[n], w2[n % 2]), n
    torch.cuda.empty_cache()
