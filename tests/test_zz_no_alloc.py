"""fn2_downsample_multi_forward_noalloc: the multi-scale downsample for callers that cannot obtain device memory.

Who that is: the ctypes stub of INTEGRATION.md section 1 when it runs inside a TensorFlow process.  TensorFlow's own
allocator has claimed the GPU's memory, the stub sees only the tensors the graph hands it, and it can neither reach
that allocator from Python nor expect a cudaMalloc next to it to succeed, so it has no `workspace` to pass to
fn2_downsample_multi_forward.  The no-alloc entry point takes inputs and outputs only and keeps its scratch in a 16 MiB
device array that is part of libfn2ops.so.

What is checked, straight through the C ABI (ctypes, raw device pointers): its outputs against those of the existing
entry point given an ordinary workspace, on two calls in a row with different shapes (the second call finds the first
one's tables and partial sums in the array and must not be affected by them), then the first shapes again; and that
shapes beyond the array are refused with a message that names the limit.  Same tolerance as test_gpu_downsample_multi.py
(1e-5 relative + 5e-6 absolute; both entry points run the same kernels in a fixed summation order, so in practice the
outputs are equal).

This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here should not hide the parity
tests.  It passes on its own in a fresh process (then its first call is also the device's first use of the array) and
after the rest of the suite.
"""
import ctypes

import pytest
import torch

from conftest import assert_close

pytestmark = pytest.mark.gpu

from flownet2_tf_b200 import _lib  # noqa: E402

BUILTIN_BYTES = 16 * 1024 * 1024   # FN2_DOWNSAMPLE_MULTI_BUILTIN_SCRATCH_BYTES (include/fn2ops.h)
LOSS_SIZES = [(6, 8), (12, 16), (24, 32), (48, 64), (96, 128)]   # flownet_c.py:135-160


def _flow(N, H, W, C, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(N, H, W, C, generator=g) * 5
    x.view(-1)[::101] = float("nan")   # the NaN-aware path as well
    return x.cuda()


def _call(x, sizes, noalloc):
    """One call of either entry point through ctypes; returns the output tensors (allocated here, by the caller)."""
    lib = _lib.lib()
    N, H, W, C = x.shape
    outs = [torch.full((N, h, w, C), -123.0, device=x.device) for h, w in sizes]
    ptrs = (ctypes.c_void_p * len(outs))(*[t.data_ptr() for t in outs])
    hw = (ctypes.c_int * (2 * len(sizes)))(*[v for s in sizes for v in s])
    stream = ctypes.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
    if noalloc:
        rc = lib.fn2_downsample_multi_forward_noalloc(x.data_ptr(), ptrs, hw, len(sizes), N, H, W, C, stream)
    else:
        nbytes = lib.fn2_downsample_multi_workspace_bytes(hw, len(sizes), N, H, W, C)
        ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=x.device)
        rc = lib.fn2_downsample_multi_forward(x.data_ptr(), ptrs, hw, len(sizes), N, H, W, C,
                                              ctypes.c_void_p(ws.data_ptr()), nbytes, stream)
    assert rc == 0, _lib.error_string(rc)
    torch.cuda.synchronize(x.device)
    return [t.cpu().numpy() for t in outs]


def test_noalloc_matches_the_workspace_entry_point_on_consecutive_calls():
    calls = [
        ((4, 384, 512, 2), LOSS_SIZES, 11),                        # the loss's call: all five scales fused
        ((1, 100, 130, 2), [(7, 9), (13, 33), (50, 40)], 12),      # other tables and a ragged last band, right after
        ((4, 384, 512, 2), LOSS_SIZES, 13),                        # the first shapes again, over the leftovers of the second
        ((2, 30, 40, 3), [(3, 4), (6, 8)], 14),                    # C = 3: single-scale kernels underneath, no scratch used
    ]
    for shape, sizes, seed in calls:
        x = _flow(*shape, seed)
        want = _call(x, sizes, noalloc=False)
        got = _call(x, sizes, noalloc=True)
        for size, g, w in zip(sizes, got, want):
            assert_close(g, w, rel=1e-5, abs_=5e-6, what="noalloc %s -> %s" % (shape, size))


def test_noalloc_refuses_shapes_beyond_the_builtin_array():
    lib = _lib.lib()
    N, H, W, C = 32, 384, 512, 2
    hw = (ctypes.c_int * (2 * len(LOSS_SIZES)))(*[v for s in LOSS_SIZES for v in s])
    assert lib.fn2_downsample_multi_workspace_bytes(hw, len(LOSS_SIZES), N, H, W, C) > BUILTIN_BYTES
    x = torch.zeros(N, H, W, C, device="cuda")
    outs = [torch.full((N, h, w, C), -123.0, device="cuda") for h, w in LOSS_SIZES]
    ptrs = (ctypes.c_void_p * len(outs))(*[t.data_ptr() for t in outs])
    rc = lib.fn2_downsample_multi_forward_noalloc(x.data_ptr(), ptrs, hw, len(LOSS_SIZES), N, H, W, C,
                                                  ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -13
    assert "16 MiB" in _lib.error_string(rc)
    torch.cuda.synchronize()
    assert all(bool((t == -123.0).all()) for t in outs)   # refused before anything was enqueued
    # the batch the header promises still fits, the next one does not
    assert lib.fn2_downsample_multi_workspace_bytes(hw, len(LOSS_SIZES), 18, H, W, C) <= BUILTIN_BYTES
    assert lib.fn2_downsample_multi_workspace_bytes(hw, len(LOSS_SIZES), 19, H, W, C) > BUILTIN_BYTES
