"""Tiny driver for ncu: runs one correlation op a few times at the bench shape.  python tools/prof_corr.py <prec> <fwd|bwd> [N]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import flownet2_tf_b200 as fn2  # noqa: E402
from conftest import corr_inputs, grad_like  # noqa: E402

prec, what = sys.argv[1], sys.argv[2]
N = int(sys.argv[3]) if len(sys.argv) > 3 else 32
a, b = corr_inputs(N, 48, 64, 256)
a, b = a.cuda(), b.cuda()
g = grad_like((N, 48, 64, 441)).cuda()
for _ in range(4):
    if what == "fwd":
        fn2.correlation(a, b, 1, 20, 1, 2, 20, precision=prec)
    else:
        fn2.correlation_grad(g, a, b, 1, 20, 1, 2, 20, precision=prec)
torch.cuda.synchronize()
print("ok")
