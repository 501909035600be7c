"""CPU tests of the boundary: libfn2ops.so loads without a GPU, exports every symbol include/fn2ops.h declares, and
its argument validation (the OP_REQUIRES of the reference OpKernels) answers before any CUDA call is made."""
import ctypes
import os
import re

import pytest
import torch

import flownet2_tf_b200 as fn2
from flownet2_tf_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "fn2ops.h")).read()
    return sorted(set(re.findall(r"FN2_API[^;(]*?\b(fn2_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = _declared_symbols()
    assert len(names) >= 11
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(handle, n), n
    assert sorted(_lib.EXPORTED_SYMBOLS) == names  # the ctypes table binds exactly the declared ABI
    assert _lib.lib().fn2_abi_version() == _lib.ABI_VERSION == 2


def test_header_compiles_as_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "fn2ops.h"\nint main(void){return FN2_ABI_VERSION - 2;}\n')
    rc = os.system("/usr/bin/gcc -std=c99 -Wall -Werror -I%s/include -c %s -o %s/t.o" % (ROOT, src, tmp_path))
    assert rc == 0


def test_shape_function_and_error_codes():
    L = _lib.lib()
    oh, ow, oc = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    assert L.fn2_correlation_output_shape(48, 64, 1, 20, 1, 2, 20, oh, ow, oc) == 0
    assert (oh.value, ow.value, oc.value) == (48, 64, 441)       # flownet_c.py:40
    assert L.fn2_correlation_output_shape(10, 12, 3, 3, 2, 2, 3, oh, ow, oc) == 0
    assert (oh.value, ow.value, oc.value) == (4, 5, 9)
    assert L.fn2_correlation_output_shape(48, 64, 2, 20, 1, 2, 20, oh, ow, oc) == -1   # even kernel
    assert L.fn2_correlation_output_shape(4, 64, 1, 20, 1, 2, 0, oh, ow, oc) == -2     # height does not fit
    assert L.fn2_correlation_output_shape(64, 4, 1, 20, 1, 2, 0, oh, ow, oc) == -3     # width does not fit
    assert L.fn2_correlation_output_shape(48, 64, 1, 20, 0, 2, 20, oh, ow, oc) == -5   # stride 0
    assert _lib.error_string(-1) == "kernel_size must be odd"                        # correlation_kernel.cc:23
    assert "fit in input height" in _lib.error_string(-2)                             # correlation_kernel.cc:50
    # validation runs before any pointer is touched or any CUDA call is made
    assert L.fn2_correlation_forward(None, None, None, 1, 48, 64, 256, 2, 20, 1, 2, 20, 0, None, 0, None) == -1
    assert L.fn2_correlation_forward(None, None, None, 1, 48, 64, 256, 1, 20, 1, 2, 20, 0, None, 0, None) == -6
    assert L.fn2_correlation_forward(16, 16, 16, 1, 48, 64, 256, 1, 20, 1, 2, 20, 9, None, 0, None) == -7
    assert L.fn2_correlation_backward(None, None, None, None, None, 1, 8, 8, 4, 1, 2, 1, 1, 2, 0, None, 0, None) == -6
    assert L.fn2_flow_warp_forward(None, None, None, 1, 8, 8, 3, None) == -6
    assert L.fn2_flow_warp_forward(8, 8, 8, 1, 8, 8, 3, None) == -6                   # misaligned
    assert L.fn2_flow_warp_backward(None, None, None, None, None, 0, 8, 8, 3, None) == -4
    assert L.fn2_downsample_forward(None, None, 1, 8, 8, 2, 1, 4, None) == -9          # division by zero case
    assert L.fn2_downsample_forward(None, None, 1, 8, 8, 2, 4, 4, None) == -6
    assert L.fn2_correlation_workspace_bytes(1, 48, 64, 256, 1, 20, 1, 2, 20, 0, 0) == 0
    # which arithmetic covers which call (no CUDA call involved)
    assert L.fn2_correlation_precision_supported(1, 48, 64, 256, 1, 20, 1, 2, 20, 0) == 1
    assert L.fn2_correlation_precision_supported(1, 48, 64, 256, 1, 20, 1, 2, 20, 1) == 1
    assert L.fn2_correlation_precision_supported(1, 48, 64, 256, 1, 20, 1, 2, 20, 3) == 1
    assert L.fn2_correlation_precision_supported(1, 48, 64, 256, 1, 20, 1, 2, 20, 2) == 0   # reserved value
    assert L.fn2_correlation_precision_supported(1, 48, 64, 256, 3, 20, 1, 2, 20, 3) == 0   # kernel_size 3
    assert L.fn2_correlation_precision_supported(1, 48, 64, 100, 1, 20, 1, 2, 20, 1) == 0   # C % 32 != 0
    assert L.fn2_correlation_precision_supported(1, 48, 64, 256, 2, 20, 1, 2, 20, 3) == -1  # the op's own error
    assert L.fn2_correlation_forward_fused(16, 16, 16, 1, 48, 64, 256, 1, 20, 1, 2, 20, 473, -0.1, 1, None, 0,
                                           None) == -5                                    # slope must be > 0


def test_python_ops_reject_what_the_reference_rejects():
    a = torch.zeros(1, 8, 8, 4)
    with pytest.raises(ValueError, match="CUDA tensor"):          # no CPU kernel exists, nor a fallback
        fn2.correlation(a, a, 1, 2, 1, 1, 2)
    with pytest.raises(ValueError, match="rank 4"):               # correlation_kernel.cc:31-32
        fn2.correlation(torch.zeros(8, 8, 4), a, 1, 2, 1, 1, 2)
    with pytest.raises(TypeError):
        fn2.correlation(a.double(), a.double(), 1, 2, 1, 1, 2)
    with pytest.raises(ValueError, match="rank 4"):               # flow_warp.cc:22-24
        fn2.flow_warp(torch.zeros(8, 8, 3), torch.zeros(1, 8, 8, 2))
    with pytest.raises(ValueError, match="2 dimensions"):         # downsample_kernel.cc:19
        fn2.downsample(a, [4, 4, 4])
    with pytest.raises(ValueError):
        fn2.correlation(a, a, 1, 2, 1, 1, 2, precision="fp64")


def test_signatures_match_the_reference_wrappers():
    import inspect
    assert list(inspect.signature(fn2.correlation).parameters)[:7] == [
        "input_a", "input_b", "kernel_size", "max_displacement", "stride_1", "stride_2", "padding"]  # correlation.py:7
    assert list(inspect.signature(fn2.flow_warp).parameters) == ["image", "flow"]                   # flow_warp.py:7
    assert list(inspect.signature(fn2.downsample).parameters) == ["tensor", "size"]                 # downsample.py:7


def test_product_does_not_import_the_oracle():
    """The product path must not route through oracle/ (or any CPU fallback)."""
    pkg = os.path.join(ROOT, "flownet2-tf_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("no oracle", ""), os.path.join(dirpath, f)
