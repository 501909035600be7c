"""Where one CTA of the fused multi-scale downsample spends its clocks (developer build: make -C flownet2-tf_b200/csrc
timeline).  python tools/timeline_ds.py [N]"""
import ctypes, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import flownet2_tf_b200 as fn2
from flownet2_tf_b200 import _lib
_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libfn2ops_tl.so")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 32
x = torch.randn(N, 384, 512, 2, device="cuda")
sizes = [(6, 8), (12, 16), (24, 32), (48, 64), (96, 128)]
for _ in range(3):
    fn2.downsample_multi(x, sizes)
torch.cuda.synchronize()
buf = np.zeros(16, np.int64)
h = _lib.lib()
h.fn2_debug_timeline_ds.argtypes = [ctypes.c_void_p]
assert h.fn2_debug_timeline_ds(buf.ctypes.data) == 0
names = ["tables", "row walk", "last rows (2 horizontal passes per scale)", "wait for the cluster", "combine", "final cluster sync"]
for i, nm in enumerate(names):
    print("%-45s %8d clk" % (nm, buf[i + 1] - buf[i]))
