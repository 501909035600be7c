"""In-kernel timeline of the CTA-pair correlation forward (developer build: make -C flownet2-tf_b200/csrc timeline).
python tools/timeline.py [N]"""
import ctypes, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import flownet2_tf_b200 as fn2
from flownet2_tf_b200 import _lib
_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libfn2ops_tl.so")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dev = torch.device("cuda", 0)
a = torch.randn(N, 48, 64, 256, device=dev); b = torch.randn(N, 48, 64, 256, device=dev)
for _ in range(3):
    fn2.correlation(a, b, 1, 20, 1, 2, 20, precision="tf32")
torch.cuda.synchronize()
buf = np.zeros(1024 * 64, np.int64)
h = _lib.lib()
h.fn2_debug_timeline.argtypes = [ctypes.c_void_p]
assert h.fn2_debug_timeline(buf.ctypes.data) == 0
t = buf.reshape(1024, 64)
g0 = t[:768, 1].min()
print("cta sm nblk | start_us end_us | sync a_full | per block: [mma_tmem_empty mma_first_full mma_commit | epi_full epi_done] (clk since entry)")
for c in list(range(0, 8)) + list(range(300, 306)) + list(range(600, 604)):
    r = t[c]; e = r[0]; nb = int(r[48])
    s = "%3d %3d %d | %7.2f %7.2f | %6d %6d |" % (c, r[47], nb, (r[1] - g0) / 1e3, (r[46] - g0) / 1e3, r[2] - e, (r[3] - e) if r[3] else -1)
    for k in range(nb):
        s += " [%d %d %d | %d %d]" % tuple((r[x + k] - e) if r[x + k] else -1 for x in (4, 12, 20, 28, 36))
    s += " | end %d %d" % (r[44] - e, r[45] - e)
    print(s)
dur = (t[:768, 46] - t[:768, 1]) / 1e3
print("CTA duration us: mean %.2f min %.2f max %.2f; kernel span %.2f us" % (dur.mean(), dur.min(), dur.max(), (t[:768, 46].max() - g0) / 1e3))
for nb in (5, 6):
    m = t[:768, 48] == nb
    print("nblk", nb, "count", int(m.sum()), "mean dur %.2f us" % dur[m].mean(), "mean clk", float((t[:768, 45] - t[:768, 0])[m].mean()))
