import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import flownet2_tf_b200 as fn2
from oracle import oracle as o
g = torch.Generator().manual_seed(1237)
x = torch.randn(2, 384, 512, 2, generator=g) * 5
nan = len(sys.argv) > 1 and sys.argv[1] == "nan"
if nan:
    x[0, 100:140, 200:300] = float("nan")
    x.view(-1)[::97] = float("nan")
for size in [(96, 128), (48, 64)]:
    got = fn2.downsample(x.cuda(), list(size)).cpu().numpy().astype(np.float64)
    want = o.c_downsample(x.numpy(), size).astype(np.float64)
    ref64 = o.np_downsample(x.numpy(), size)
    for name, A, B in [("gpu-vs-c32", got, want), ("gpu-vs-np64", got, ref64), ("c32-vs-np64", want, ref64)]:
        err = np.abs(np.nan_to_num(A - B))
        print(size, name, "max err", err.max(), "n>1e-5:", int((err > 1e-5).sum()), "argmax", np.unravel_index(err.argmax(), err.shape))
size = (96, 128)
got = fn2.downsample(x.cuda(), list(size)).cpu().numpy().astype(np.float64)
ref64 = o.np_downsample(x.numpy(), size)
err = np.abs(np.nan_to_num(got - ref64))
np.set_printoptions(linewidth=250, precision=1)
print("per-dy max err x1e6:", err.max(axis=(0, 2, 3)) * 1e6)
print("per-dx max err x1e6:", err.max(axis=(0, 1, 3)) * 1e6)
