"""correlation(input_a, input_b, kernel_size, max_displacement, stride_1, stride_2, padding)

Drop-in for the reference's ``src/correlation.py:7-14`` (forward) and ``:17-35`` (registered gradient), on NHWC
float32 CUDA tensors.  Shape and error behaviour follow ``src/ops/correlation/correlation_op.cc:9-62`` and
``correlation_kernel.cc:23-53``.
"""
import ctypes
import os
import warnings

import torch

from . import _lib

#: Arithmetic of the contraction when the caller does not pass ``precision=``.  "auto" (default) is the fastest mode
#: that meets the reference's float32 tolerance (1e-5 rel / 1e-6 abs): "tf32x3" (split-tf32 on the tensor cores) where
#: the call has the FlowNetC shape, else "fp32" (CUDA cores, any geometry).  "tf32" is faster and less accurate
#: (2.5e-3 of the tensor's maximum).  Override with FN2_CORR_PRECISION.
DEFAULT_PRECISION = os.environ.get("FN2_CORR_PRECISION", "auto")

_NAMES = {v: n for n, v in _lib.PRECISIONS.items()}


def _prec(precision, geom):
    """Resolve ``precision`` for one call (geom = N,H,W,C,k,md,s1,s2,pad) to the ABI's enum value.  A tensor-core
    mode asked for on a geometry it does not cover falls back to "fp32" with a warning (never to less accuracy)."""
    p = DEFAULT_PRECISION if precision is None else precision
    if p != "auto" and p not in _lib.PRECISIONS:
        raise ValueError("precision must be 'auto' or one of %s" % sorted(_lib.PRECISIONS))
    want = _lib.PRECISIONS["tf32x3" if p == "auto" else p]
    if want == 0:
        return 0
    ok = _lib.lib().fn2_correlation_precision_supported(*geom, want)
    if ok < 0:
        _lib.check(ok, "Correlation")
    if ok == 1:
        return want
    if p != "auto":
        warnings.warn("correlation: precision=%r does not cover this call (needs kernel_size 1, stride_1 1, "
                      "pad == max_displacement, max_displacement // stride_2 == 10, C %% 32 == 0, C <= 256); "
                      "using the fp32 CUDA-core kernels" % p, RuntimeWarning, stacklevel=3)
    return 0


def output_shape(H, W, kernel_size, max_displacement, stride_1, stride_2, padding):
    """[out_h, out_w, out_channels] -- the op's shape function (correlation_op.cc:9-51)."""
    oh, ow, oc = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    st = _lib.lib().fn2_correlation_output_shape(H, W, kernel_size, max_displacement, stride_1, stride_2, padding,
                                                 ctypes.byref(oh), ctypes.byref(ow), ctypes.byref(oc))
    _lib.check(st, "Correlation")
    return oh.value, ow.value, oc.value


def _check_pair(a, b):
    _lib.require_cuda_f32(a, "input_a")
    _lib.require_cuda_f32(b, "input_b")
    if a.shape != b.shape:  # shape fn merges the two input shapes (correlation_op.cc:17)
        raise ValueError("input_a and input_b must have the same shape, got %s and %s"
                         % (tuple(a.shape), tuple(b.shape)))
    if a.device != b.device:
        raise ValueError("input_a and input_b must be on the same device")


def _workspace(a, geom, prec, direction):
    nbytes = _lib.lib().fn2_correlation_workspace_bytes(*geom, prec, direction)
    if nbytes == 0:
        return None, 0, ctypes.c_void_p(0)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=a.device)
    return ws, nbytes, ctypes.c_void_p(ws.data_ptr())


def _forward(a, b, k, md, s1, s2, pad, prec):
    _check_pair(a, b)
    N, H, W, C = a.shape
    oh, ow, oc = output_shape(H, W, k, md, s1, s2, pad)
    out = torch.empty((N, oh, ow, oc), dtype=torch.float32, device=a.device)
    geom = (N, H, W, C, k, md, s1, s2, pad)
    with torch.cuda.device(a.device):
        ws, nbytes, wsp = _workspace(a, geom, prec, 0)
        st = _lib.lib().fn2_correlation_forward(a.data_ptr(), b.data_ptr(), out.data_ptr(), *geom, prec, wsp, nbytes,
                                                _lib.current_stream_ptr(a.device))
    _lib.check(st, "Correlation")
    return out


def correlation_grad(gradients, input_a, input_b, kernel_size, max_displacement, stride_1, stride_2, pad,
                     precision=None):
    """The ``CorrelationGrad`` op (correlation_op.cc:64-82): returns (backprops_a, backprops_b)."""
    _check_pair(input_a, input_b)
    _lib.require_cuda_f32(gradients, "gradients")
    N, H, W, C = input_a.shape
    prec = precision if isinstance(precision, int) else _prec(
        precision, (N, H, W, C, kernel_size, max_displacement, stride_1, stride_2, pad))
    oh, ow, oc = output_shape(H, W, kernel_size, max_displacement, stride_1, stride_2, pad)
    if tuple(gradients.shape) != (N, oh, ow, oc):
        raise ValueError("gradients must have shape %s, got %s" % ((N, oh, ow, oc), tuple(gradients.shape)))
    ga = torch.empty_like(input_a)
    gb = torch.empty_like(input_b)
    geom = (N, H, W, C, kernel_size, max_displacement, stride_1, stride_2, pad)
    with torch.cuda.device(input_a.device):
        ws, nbytes, wsp = _workspace(input_a, geom, prec, 1)
        st = _lib.lib().fn2_correlation_backward(gradients.data_ptr(), input_a.data_ptr(), input_b.data_ptr(),
                                                 ga.data_ptr(), gb.data_ptr(), *geom, prec, wsp, nbytes,
                                                 _lib.current_stream_ptr(input_a.device))
    _lib.check(st, "CorrelationGrad")
    return ga, gb


class _Correlation(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, k, md, s1, s2, pad, prec):
        ctx.save_for_backward(a, b)
        ctx.attrs = (k, md, s1, s2, pad, prec)
        return _forward(a, b, k, md, s1, s2, pad, prec)

    @staticmethod
    def backward(ctx, gradients):
        a, b = ctx.saved_tensors
        k, md, s1, s2, pad, prec = ctx.attrs
        ga, gb = correlation_grad(gradients.contiguous(), a, b, k, md, s1, s2, pad, precision=prec)
        return ga, gb, None, None, None, None, None, None


def correlation(input_a, input_b, kernel_size, max_displacement, stride_1, stride_2, padding, *, precision=None):
    """Cost volume between two feature maps; same positional signature as the reference wrapper.

    ``precision`` is the one keyword-only extension: "auto" (default: "tf32x3" where the tensor-core kernels cover the
    call, else "fp32" -- both meet the reference's float32 tolerance), "fp32", "tf32x3" or "tf32".
    """
    args = [int(v) for v in (kernel_size, max_displacement, stride_1, stride_2, padding)]
    _check_pair(input_a, input_b)
    return _Correlation.apply(input_a, input_b, *args, _prec(precision, tuple(input_a.shape) + tuple(args)))


# ---- fused forward epilogue ----------------------------------------------------------
class _CorrelationLeakyConcat(torch.autograd.Function):
    """cc_relu = LeakyReLU(correlation(a, b, ...)) written straight into channels [offset, offset + D) of ``buf``.

    Backward: d cc = d buf[..., offset:offset+D] * (cc_relu > 0 ? 1 : slope), formed inside the CorrelationGrad
    kernel's gradient gather (``fn2_correlation_backward_fused``); the other channels of d buf pass through to
    ``buf``'s producer."""

    @staticmethod
    def forward(ctx, a, b, buf, offset, slope, k, md, s1, s2, pad, prec):
        _check_pair(a, b)
        _lib.require_cuda_f32(buf, "concat buffer")
        N, H, W, C = a.shape
        oh, ow, oc = output_shape(H, W, k, md, s1, s2, pad)
        if tuple(buf.shape[:3]) != (N, oh, ow) or offset < 0 or offset + oc > buf.shape[3]:
            raise ValueError("concat buffer %s cannot hold [%d,%d,%d,%d] at channel %d"
                             % (tuple(buf.shape), N, oh, ow, oc, offset))
        if offset % 4 != 0:
            raise ValueError("channel offset must be a multiple of 4 (16-byte alignment of the first element)")
        geom = (N, H, W, C, k, md, s1, s2, pad)
        with torch.cuda.device(a.device):
            ws, nbytes, wsp = _workspace(a, geom, prec, 0)
            st = _lib.lib().fn2_correlation_forward_fused(
                a.data_ptr(), b.data_ptr(), buf.data_ptr() + 4 * offset, *geom, int(buf.shape[3]), float(slope), prec,
                wsp, nbytes, _lib.current_stream_ptr(a.device))
        _lib.check(st, "Correlation")
        ctx.mark_dirty(buf)
        ctx.save_for_backward(a, b, buf)
        ctx.attrs = (offset, oc, slope, k, md, s1, s2, pad, prec)
        return buf

    @staticmethod
    def backward(ctx, gbuf):
        a, b, buf = ctx.saved_tensors
        offset, oc, slope, k, md, s1, s2, pad, prec = ctx.attrs
        gbuf = gbuf.contiguous()
        N, H, W, C = a.shape
        geom = (N, H, W, C, k, md, s1, s2, pad)
        ga, gb = torch.empty_like(a), torch.empty_like(b)
        # slice + LeakyReLU derivative inside the CorrelationGrad kernel's gradient gather: no masked copy of the
        # 441 channels is made (fn2_correlation_backward_fused)
        with torch.cuda.device(a.device):
            ws, nbytes, wsp = _workspace(a, geom, prec, 1)
            st = _lib.lib().fn2_correlation_backward_fused(
                gbuf.data_ptr() + 4 * offset, buf.data_ptr() + 4 * offset, a.data_ptr(), b.data_ptr(), ga.data_ptr(),
                gb.data_ptr(), *geom, int(buf.shape[3]), float(slope), prec, wsp, nbytes,
                _lib.current_stream_ptr(a.device))
        _lib.check(st, "CorrelationGrad")
        # what flows on to the buffer's producer: the channels this op did not overwrite (conv_redir's 32)
        gpass = torch.zeros_like(gbuf)
        gpass[..., :offset] = gbuf[..., :offset]
        gpass[..., offset + oc:] = gbuf[..., offset + oc:]
        return ga, gb, gpass, None, None, None, None, None, None, None, None


def correlation_leaky_concat(input_a, input_b, kernel_size, max_displacement, stride_1, stride_2, padding, buf,
                             channel_offset, leaky_slope=0.1, *, precision=None):
    """LeakyReLU(correlation(...)) stored into ``buf[..., channel_offset : channel_offset + D]`` by the kernel's own
    epilogue; returns ``buf``.  Replaces the three graph nodes of ``flownet_c.py:40-46`` (correlation, LeakyReLU,
    tf.concat) -- the [N,h,w,441] volume makes no round trip through HBM.  ``buf`` is a contiguous NHWC float32
    tensor whose other channels (conv_redir's 32) the caller fills, before or after."""
    args = [int(v) for v in (kernel_size, max_displacement, stride_1, stride_2, padding)]
    _check_pair(input_a, input_b)
    if not float(leaky_slope) > 0:   # the backward reads the mask off the sign of the stored output
        raise ValueError("leaky_slope must be > 0, got %r" % (leaky_slope,))
    return _CorrelationLeakyConcat.apply(input_a, input_b, buf, int(channel_offset), float(leaky_slope), *args,
                                         _prec(precision, tuple(input_a.shape) + tuple(args)))
