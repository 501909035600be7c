"""Loader for oracle/_ref/libfn2ref.so: the reference's OWN CUDA kernels (src/ops/**.cu.cc, compiled unmodified by
oracle/Makefile with the header stand-ins in oracle/tf_stub).  TEST INFRASTRUCTURE ONLY.  Needs a GPU.

Used to pin the oracle to the real reference (tests/test_reference_pin.py), to generate tests/golden/*.npz
(tests/golden/make_golden.py), and to time the reference's GPU kernels beside ours (bench.py, informational).
"""
import ctypes
import os

import torch

_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libfn2ref.so")
_lib = None
_p, _i = ctypes.c_void_p, ctypes.c_int


def available():
    return os.path.exists(_SO)


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(_SO)
        L.fn2ref_correlation_forward.argtypes = [_p] * 3 + [_i] * 9 + [_p] * 2
        L.fn2ref_correlation_backward.argtypes = [_p] * 5 + [_i] * 9 + [_p] * 2
        L.fn2ref_flow_warp_forward.argtypes = [_p] * 3 + [_i] * 4
        L.fn2ref_flow_warp_backward.argtypes = [_p] * 5 + [_i] * 4
        L.fn2ref_downsample_forward.argtypes = [_p] * 2 + [_i] * 6
        _lib = L
    return _lib


def _sync_in():
    # the reference launches on the legacy default stream; make torch's stream work visible to it and back
    torch.cuda.synchronize()


def correlation(a, b, k, md, s1, s2, pad, out=None, temps=None):
    N, H, W, C = a.shape
    from . import oracle as _o
    oh, ow, oc = _o.out_shape(H, W, k, md, s1, s2, pad)
    out = torch.empty((N, oh, ow, oc), device=a.device) if out is None else out
    if temps is None:
        temps = (torch.empty((N, H + 2 * pad, W + 2 * pad, C), device=a.device),
                 torch.empty((N, H + 2 * pad, W + 2 * pad, C), device=a.device))
    _sync_in()
    rc = lib().fn2ref_correlation_forward(a.data_ptr(), b.data_ptr(), out.data_ptr(), N, H, W, C, k, md, s1, s2, pad,
                                          temps[0].data_ptr(), temps[1].data_ptr())
    assert rc == 0, rc
    return out


def correlation_grad(g, a, b, k, md, s1, s2, pad, temps=None):
    N, H, W, C = a.shape
    ga, gb = torch.empty_like(a), torch.empty_like(b)
    if temps is None:
        temps = (torch.empty((N, H + 2 * pad, W + 2 * pad, C), device=a.device),
                 torch.empty((N, H + 2 * pad, W + 2 * pad, C), device=a.device))
    _sync_in()
    rc = lib().fn2ref_correlation_backward(g.data_ptr(), a.data_ptr(), b.data_ptr(), ga.data_ptr(), gb.data_ptr(),
                                           N, H, W, C, k, md, s1, s2, pad, temps[0].data_ptr(), temps[1].data_ptr())
    assert rc == 0, rc
    return ga, gb


def flow_warp(image, flow):
    N, H, W, C = image.shape
    out = torch.empty_like(image)
    _sync_in()
    assert lib().fn2ref_flow_warp_forward(image.data_ptr(), flow.data_ptr(), out.data_ptr(), N, H, W, C) == 0
    return out


def flow_warp_grad(image, flow, g):
    N, H, W, C = image.shape
    gi, gf = torch.empty_like(image), torch.empty_like(flow)
    _sync_in()
    assert lib().fn2ref_flow_warp_backward(image.data_ptr(), flow.data_ptr(), g.data_ptr(), gi.data_ptr(),
                                           gf.data_ptr(), N, H, W, C) == 0
    return gi, gf


def downsample(x, size):
    N, H, W, C = x.shape
    out = torch.empty((N, size[0], size[1], C), device=x.device)
    _sync_in()
    assert lib().fn2ref_downsample_forward(x.data_ptr(), out.data_ptr(), N, H, W, C, size[0], size[1]) == 0
    return out
