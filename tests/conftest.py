import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


# ---- shared synthetic inputs ----------------------------------------------------------
def corr_inputs(N, H, W, C, seed=1234):
    import torch
    g = torch.Generator().manual_seed(seed)
    a = torch.nn.functional.leaky_relu(torch.randn(N, H, W, C, generator=g), 0.1)
    b = torch.nn.functional.leaky_relu(torch.randn(N, H, W, C, generator=g), 0.1)
    return a.contiguous(), b.contiguous()


def grad_like(shape, seed=1235):
    import torch
    g = torch.Generator().manual_seed(seed)
    return torch.randn(*shape, generator=g)


def warp_inputs(N, H, W, C=3, seed=1236):
    """image in [0,1); flow ~ 8 px with 5 % far out of range, 1 % exactly integer, a few NaN."""
    import torch
    g = torch.Generator().manual_seed(seed)
    image = torch.rand(N, H, W, C, generator=g)
    flow = torch.randn(N, H, W, 2, generator=g) * 8
    u = torch.rand(N, H, W, generator=g)
    far = u < 0.05
    flow[far] = torch.where(torch.rand(int(far.sum()), 2, generator=g) < 0.5, -2.0 * W, 2.0 * W)
    integer = (u >= 0.05) & (u < 0.06)
    flow[integer] = flow[integer].round()
    flow.view(-1, 2)[::max(1, (N * H * W) // 7)] = float("nan")
    return image.contiguous(), flow.contiguous()


def assert_close(got, want, rel=1e-5, abs_=1e-6, what=""):
    """|got - want| <= abs_ + rel * |want| elementwise (the tolerance BASELINE.json's north_star states)."""
    got = np.asarray(got, np.float64)
    want = np.asarray(want, np.float64)
    assert got.shape == want.shape, (got.shape, want.shape)
    nan_g, nan_w = np.isnan(got), np.isnan(want)
    assert (nan_g == nan_w).all(), "%s: NaN pattern differs" % what
    err = np.abs(np.where(nan_w, 0, got - want))
    lim = abs_ + rel * np.abs(np.where(nan_w, 0, want))
    bad = err > lim
    assert not bad.any(), "%s: %d/%d outside tolerance, max err %.3e (limit there %.3e)" % (
        what, int(bad.sum()), bad.size, float(err.max()), float(lim.flat[int(err.argmax())]))


# ---- oracle for the row-halo (strip) variants: the full op on zero-extended tensors, cropped -------------------
def oracle_halo_forward(a, b_ext, k, md, s1, s2, pad, top, bottom, precision=None):
    """numpy/torch-CPU restatement of fn2_correlation_forward_halo: extend a with zero rows to b_ext's height, run
    the whole-image oracle, keep the strip's rows."""
    import torch
    from oracle import oracle as o
    a_np = a.detach().cpu().numpy()
    N, H, W, C = a_np.shape
    a_ext = np.zeros((N, H + top + bottom, W, C), np.float32)
    a_ext[:, top:top + H] = a_np
    full = o.c_correlation(a_ext, np.ascontiguousarray(b_ext.detach().cpu().numpy()), k, md, s1, s2, pad, f64=True)
    return torch.from_numpy(np.ascontiguousarray(full[:, top:top + H]).astype(np.float32))


def oracle_halo_backward(g, a, b_ext, k, md, s1, s2, pad, top, bottom, precision=None):
    import torch
    from oracle import oracle as o
    a_np, g_np = a.detach().cpu().numpy(), g.detach().cpu().numpy()
    N, H, W, C = a_np.shape
    a_ext = np.zeros((N, H + top + bottom, W, C), np.float32)
    a_ext[:, top:top + H] = a_np
    g_ext = np.zeros((N, H + top + bottom) + g_np.shape[2:], np.float32)
    g_ext[:, top:top + H] = g_np
    ga, gb = o.c_correlation_grad(g_ext, a_ext, np.ascontiguousarray(b_ext.detach().cpu().numpy()), k, md, s1, s2,
                                  pad, f64=True)
    return (torch.from_numpy(np.ascontiguousarray(ga[:, top:top + H]).astype(np.float32)),
            torch.from_numpy(gb.astype(np.float32)))


# ---- the CPU oracle behind the reference's three op signatures, for running the caller graphs without the kernels ---
def oracle_ops():
    """correlation / flow_warp / downsample as torch autograd functions over oracle/ (C restatement, float64 inside).
    Test infrastructure only: lets the graphs of flownet2_tf_b200.graphs run on CPU with the same weights as on the GPU."""
    import torch
    from types import SimpleNamespace
    from oracle import oracle as o

    def t(x):
        return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))

    class Corr(torch.autograd.Function):
        @staticmethod
        def forward(ctx, a, b, args):
            ctx.save_for_backward(a, b)
            ctx.args = args
            return t(o.c_correlation(a.detach().contiguous().numpy(), b.detach().contiguous().numpy(), *args, f64=True))

        @staticmethod
        def backward(ctx, g):
            a, b = ctx.saved_tensors
            ga, gb = o.c_correlation_grad(g.contiguous().numpy(), a.detach().contiguous().numpy(),
                                          b.detach().contiguous().numpy(), *ctx.args, f64=True)
            return t(ga), t(gb), None

    class Warp(torch.autograd.Function):
        @staticmethod
        def forward(ctx, image, flow):
            ctx.save_for_backward(image, flow)
            return t(o.c_flow_warp(image.detach().contiguous().numpy(), flow.detach().contiguous().numpy()))

        @staticmethod
        def backward(ctx, g):
            image, flow = ctx.saved_tensors
            gi, gf = o.c_flow_warp_grad(image.detach().contiguous().numpy(), flow.detach().contiguous().numpy(),
                                        g.contiguous().numpy())
            return t(gi), t(gf)

    def downsample(x, size):
        return t(o.c_downsample(x.detach().contiguous().numpy(), tuple(int(v) for v in size)))

    return SimpleNamespace(
        correlation=lambda a, b, k, md, s1, s2, pad: Corr.apply(a, b, (k, md, s1, s2, pad)),
        flow_warp=Warp.apply, downsample=downsample,
        downsample_multi=lambda x, sizes: [downsample(x, s) for s in sizes])
