"""downsample(tensor, size) -- drop-in for the reference's ``src/downsample.py:7-8``.

``size`` = [out_height, out_width] (the op attr, ``src/ops/downsample/downsample_op.cc:24-28``).  The reference
registers no gradient for this op; differentiating through it raises LookupError here as tf.gradients does there.
"""
import torch

from . import _lib


def _forward(tensor, size):
    _lib.require_cuda_f32(tensor, "input")  # "Input images must have rank 4" (downsample_kernel.cc:25)
    size = list(size)
    if len(size) != 2:  # downsample_kernel.cc:19
        raise ValueError("size must be 2 dimensions")
    oh, ow = int(size[0]), int(size[1])
    N, H, W, C = tensor.shape
    if oh < 1 or ow < 1:
        raise ValueError("Downsample: output dimensions must be positive")
    out = torch.empty((N, oh, ow, C), dtype=torch.float32, device=tensor.device)
    with torch.cuda.device(tensor.device):
        st = _lib.lib().fn2_downsample_forward(tensor.data_ptr(), out.data_ptr(), N, H, W, C, oh, ow,
                                               _lib.current_stream_ptr(tensor.device))
    _lib.check(st, "Downsample")
    return out


class _Downsample(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tensor, oh, ow):
        return _forward(tensor, (oh, ow))

    @staticmethod
    def backward(ctx, gradients):
        raise LookupError("No gradient defined for operation 'Downsample' (the reference registers none)")


def downsample(tensor, size):
    size = list(size)
    if len(size) != 2:
        raise ValueError("size must be 2 dimensions")
    if isinstance(tensor, torch.Tensor) and tensor.requires_grad:
        return _Downsample.apply(tensor, int(size[0]), int(size[1]))
    return _forward(tensor, size)


def downsample_multi(tensor, sizes):
    """[downsample(tensor, size) for size in sizes] in one pass over ``tensor``.

    The reference's losses downsample the same ground-truth flow to five sizes with five op calls
    (``flownet_c.py:135-160``); each output here equals the single-size op's.  No gradient, as for ``downsample``.
    """
    import ctypes
    _lib.require_cuda_f32(tensor, "input")
    if isinstance(tensor, torch.Tensor) and tensor.requires_grad:
        raise LookupError("No gradient defined for operation 'Downsample' (the reference registers none)")
    sizes = [list(s) for s in sizes]
    if not sizes:
        return []
    for s in sizes:
        if len(s) != 2:  # downsample_kernel.cc:19
            raise ValueError("size must be 2 dimensions")
        if int(s[0]) < 1 or int(s[1]) < 1:
            raise ValueError("Downsample: output dimensions must be positive")
    N, H, W, C = tensor.shape
    outs = [torch.empty((N, int(s[0]), int(s[1]), C), dtype=torch.float32, device=tensor.device) for s in sizes]
    ptrs = (ctypes.c_void_p * len(outs))(*[o.data_ptr() for o in outs])
    hw = (ctypes.c_int * (2 * len(outs)))(*[int(v) for s in sizes for v in s])
    with torch.cuda.device(tensor.device):
        st = _lib.lib().fn2_downsample_multi_forward(tensor.data_ptr(), ptrs, hw, len(outs), N, H, W, C,
                                                     _lib.current_stream_ptr(tensor.device))
    _lib.check(st, "Downsample")
    return outs
