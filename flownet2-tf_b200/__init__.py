"""B200-native drop-in for the custom ops of fperezgamonal/flownet2-tf.

The three public functions keep the reference's signatures (``src/correlation.py:7``, ``src/flow_warp.py:7``,
``src/downsample.py:7``) and operate on NHWC float32 CUDA tensors::

    from flownet2_tf_b200 import correlation, flow_warp, downsample

They call hand-written sm_100a kernels in ``libfn2ops.so`` through a C ABI (``include/fn2ops.h``) with ctypes.
There is no CPU fallback: importing works anywhere, but calling an op without the built library or without a CUDA
tensor raises.
"""
from .correlation import correlation, correlation_grad, correlation_leaky_concat  # noqa: F401
from .flow_warp import flow_warp, flow_warp_grad, flow_warp_error  # noqa: F401
from .downsample import downsample, downsample_multi, multiscale_epe  # noqa: F401
from .flo import read_flo, write_flo  # noqa: F401
from . import _lib  # noqa: F401

__all__ = ["correlation", "correlation_grad", "flow_warp", "flow_warp_grad", "downsample",
           "downsample_multi", "multiscale_epe", "correlation_leaky_concat", "flow_warp_error", "read_flo", "write_flo"]
