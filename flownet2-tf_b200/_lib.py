"""ctypes binding of libfn2ops.so (C ABI declared in include/fn2ops.h).

This is the analogue of the reference's ``tf.load_op_library("./ops/build/<name>.so")`` lines
(src/correlation.py:3-4, src/flow_warp.py:3-4, src/downsample.py:3-4).  The library is built in-tree by
``make -C flownet2-tf_b200/csrc`` (or ``__graft_entry__.build()``).  Missing library => RuntimeError: no fallback.
"""
import ctypes
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
# FN2_OPS_LIBRARY: developer override, to time a variant build of the same ABI (never a fallback: the file must exist)
LIB_PATH = os.environ.get("FN2_OPS_LIBRARY") or os.path.join(_HERE, "libfn2ops.so")
CSRC_DIR = os.path.join(_HERE, "csrc")

ABI_VERSION = 2   # FN2_ABI_VERSION of include/fn2ops.h this table was written against

PRECISIONS = {"fp32": 0, "tf32": 1, "tf32x3": 3}

_lock = threading.Lock()
_lib = None

c_int, c_size_t, c_void_p = ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p

_CORR_GEOM = [c_int] * 9  # N,H,W,C,k,md,s1,s2,pad

_SIGNATURES = {
    "fn2_abi_version": (c_int, []),
    "fn2_error_string": (ctypes.c_char_p, [c_int]),
    "fn2_last_cuda_error": (c_int, []),
    "fn2_launch_count": (ctypes.c_ulonglong, []),
    "fn2_correlation_output_shape": (c_int, [c_int] * 7 + [ctypes.POINTER(c_int)] * 3),
    "fn2_correlation_precision_supported": (c_int, _CORR_GEOM + [c_int]),
    "fn2_correlation_workspace_bytes": (c_size_t, _CORR_GEOM + [c_int, c_int]),
    "fn2_correlation_forward": (c_int, [c_void_p] * 3 + _CORR_GEOM + [c_int, c_void_p, c_size_t, c_void_p]),
    "fn2_correlation_forward_fused": (c_int, [c_void_p] * 3 + _CORR_GEOM + [c_int, ctypes.c_float, c_int, c_void_p,
                                              c_size_t, c_void_p]),
    "fn2_correlation_backward_fused": (c_int, [c_void_p] * 6 + _CORR_GEOM + [c_int, ctypes.c_float, c_int, c_void_p,
                                               c_size_t, c_void_p]),
    "fn2_correlation_backward": (c_int, [c_void_p] * 5 + _CORR_GEOM + [c_int, c_void_p, c_size_t, c_void_p]),
    "fn2_correlation_forward_halo": (c_int, [c_void_p] * 3 + _CORR_GEOM + [c_int] * 3 + [c_void_p, c_size_t, c_void_p]),
    "fn2_correlation_backward_halo": (c_int, [c_void_p] * 5 + _CORR_GEOM + [c_int] * 3 + [c_void_p, c_size_t, c_void_p]),
    "fn2_flow_warp_forward": (c_int, [c_void_p] * 3 + [c_int] * 4 + [c_void_p]),
    "fn2_flow_warp_error_forward": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p]),
    "fn2_flow_warp_backward": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p]),
    "fn2_downsample_forward": (c_int, [c_void_p] * 2 + [c_int] * 6 + [c_void_p]),
    "fn2_downsample_multi_forward": (c_int, [c_void_p, ctypes.POINTER(c_void_p), ctypes.POINTER(c_int)] + [c_int] * 5
                                     + [c_void_p, c_size_t, c_void_p]),
    "fn2_downsample_multi_forward_noalloc": (c_int, [c_void_p, ctypes.POINTER(c_void_p), ctypes.POINTER(c_int)]
                                             + [c_int] * 5 + [c_void_p]),
    "fn2_downsample_multi_workspace_bytes": (c_size_t, [ctypes.POINTER(c_int)] + [c_int] * 5),
    "fn2_multiscale_epe_edges_forward": (c_int, [c_void_p, ctypes.c_float, ctypes.POINTER(c_void_p),
                                                 ctypes.POINTER(c_void_p), ctypes.c_float, ctypes.POINTER(c_int)]
                                         + [c_int] * 5 + [ctypes.POINTER(c_void_p), ctypes.POINTER(c_void_p), c_void_p,
                                                          c_void_p, c_size_t, c_void_p]),
    "fn2_multiscale_epe_supported": (c_int, [ctypes.POINTER(c_int)] + [c_int] * 5),
    "fn2_multiscale_epe_forward": (c_int, [c_void_p, ctypes.c_float, ctypes.POINTER(c_void_p), ctypes.POINTER(c_int)]
                                   + [c_int] * 5 + [ctypes.POINTER(c_void_p), ctypes.POINTER(c_void_p), c_void_p,
                                                    c_void_p, c_size_t, c_void_p]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def build(verbose=False):
    """Compile libfn2ops.so for sm_100a (nvcc cross-compiles without a GPU)."""
    res = subprocess.run(["make", "-C", CSRC_DIR, "-j8"], capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:])
        print(res.stderr[-4000:])
    if res.returncode != 0:
        raise RuntimeError("building libfn2ops.so failed")
    return LIB_PATH


def lib():
    """Load (once) and return the ctypes handle.  Raises RuntimeError when the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    "libfn2ops.so not found at %s: build it with `make -C %s` (there is no CPU fallback)"
                    % (LIB_PATH, CSRC_DIR))
            handle = ctypes.CDLL(LIB_PATH)
            for name, (restype, argtypes) in _SIGNATURES.items():
                fn = getattr(handle, name)
                fn.restype, fn.argtypes = restype, argtypes
            if handle.fn2_abi_version() != ABI_VERSION:
                raise RuntimeError("libfn2ops.so has ABI version %d, this binding was written for %d: rebuild it "
                                   "(make -C %s)" % (handle.fn2_abi_version(), ABI_VERSION, CSRC_DIR))
            _lib = handle
    return _lib


def error_string(status):
    return lib().fn2_error_string(int(status)).decode()


def check(status, what):
    """Raise like the reference op would: InvalidArgument conditions -> ValueError, CUDA failures -> RuntimeError."""
    if status == 0:
        return
    msg = "%s: %s" % (what, error_string(status))
    if status == -10:
        raise RuntimeError(msg + " (cudaError %d)" % lib().fn2_last_cuda_error())
    raise ValueError(msg)


def launch_count():
    return int(lib().fn2_launch_count())


def require_cuda_f32(t, name, rank=4):
    """The ops are float32, NHWC, device-resident only (correlation_op.cc:54-61 etc.); anything else is rejected
    instead of being silently converted."""
    import torch
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor" % name)
    if t.dim() != rank:
        raise ValueError("%s must have rank %d" % (name, rank))
    if t.dtype != torch.float32:
        raise TypeError("%s must be float32" % name)
    if not t.is_cuda:
        raise ValueError("%s must be a CUDA tensor (there is no CPU implementation of this op)" % name)
    if not t.is_contiguous():
        raise ValueError("%s must be contiguous NHWC" % name)
    return t


def current_stream_ptr(device):
    import torch
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)
