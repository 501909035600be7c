"""GPU parity tests of the tensor-core (tcgen05) correlation paths against the f64 oracle.

Stated tolerances (BASELINE.json north_star: "the tf32/bf16 paths within a stated tolerance"):
  tf32   : |err| <= 2.5e-3 * max|ref|   (10-bit mantissa operands, fp32 accumulate over C = 256 products; measured
           worst case on these inputs is ~9e-4 * max|ref|, profiles/r01_b_check_tc_bwd.log)
  tf32x3 : the fp32 tolerance of the op, 1e-5 relative / 1e-6 absolute (3xTF32 split, fp32-grade)
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


def test_unsupported_precision_is_an_error_not_a_fallback():
    a = torch.zeros(1, 8, 8, 4, device="cuda")   # C % 32 != 0: no tensor-core tiling
    with pytest.raises(ValueError, match="precision not supported"):
        fn2.correlation(a, a, 1, 2, 1, 2, 2, precision="tf32")


@pytest.mark.parametrize("shape", [
    (40, 47, 61, 64),    # more pair tiles than resident pairs, classes of different sizes (invalid tiles in the list),
                         # partial column tiles, a tail that is split
    (75, 24, 32, 32),    # 300 pair tiles on 74 pairs: 4 rounds + a split tail of 4 tiles
    (37, 16, 32, 128),   # exactly two rounds, no tail (C a power of two: the 1/C scaling stays exact)
    (3, 48, 64, 256),    # fewer pair tiles than resident pairs (one tile each)
])
def test_persistent_pair_forward_is_exact_on_integer_inputs(shape):
    """The persistent CTA-pair forward walks a work list with running barrier phases; integer inputs make every sum
    exact, so any tile taken twice, skipped, or computed against a stale a tile shows as a wrong integer."""
    N, H, W, C = shape
    g = torch.Generator().manual_seed(sum(shape))
    a = torch.randint(-3, 4, shape, generator=g).float()
    b = torch.randint(-3, 4, shape, generator=g).float()
    out = fn2.correlation(a.cuda(), b.cuda(), *ARGS, precision="tf32")
    want = o.c_correlation(a.numpy(), b.numpy(), *ARGS, f64=True)
    np.testing.assert_array_equal(out.cpu().numpy().astype(np.float64), want)
    out2 = fn2.correlation(a.cuda(), b.cuda(), *ARGS, precision="tf32")
    assert torch.equal(out, out2)


@pytest.mark.parametrize("shape", [
    (14, 47, 61, 64),    # 336 tiles on 148 CTAs: every CTA walks over several tiles, ragged classes, partial tiles
    (37, 16, 32, 128),   # exactly two tiles per CTA, one row tile per class
    (6, 20, 140, 32),    # wide classes: three column tiles, two 32-pixel items per source row
    (3, 48, 64, 256),    # fewer tiles than CTAs (one tile each)
])
def test_persistent_backward_is_exact_on_integer_inputs(shape):
    """The persistent backward kernel carries its stage ring, row barriers and accumulator from tile to tile; integer
    inputs make every sum exact (C is a power of two), so a stale G block, an undrained accumulator row or a tile
    taken twice shows as a wrong number."""
    N, H, W, C = shape
    g = torch.Generator().manual_seed(sum(shape))
    a = torch.randint(-3, 4, shape, generator=g).float()
    b = torch.randint(-3, 4, shape, generator=g).float()
    gr = torch.randint(-3, 4, (N, H, W, 441), generator=g).float()
    ga, gb = fn2.correlation_grad(gr.cuda(), a.cuda(), b.cuda(), *ARGS, precision="tf32")
    wa, wb = o.c_correlation_grad(gr.numpy(), a.numpy(), b.numpy(), *ARGS, f64=True)
    np.testing.assert_array_equal(ga.cpu().numpy().astype(np.float64), wa)
    np.testing.assert_array_equal(gb.cpu().numpy().astype(np.float64), wb)
    ga2, gb2 = fn2.correlation_grad(gr.cuda(), a.cuda(), b.cuda(), *ARGS, precision="tf32")
    assert torch.equal(ga, ga2) and torch.equal(gb, gb2)
