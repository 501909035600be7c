"""Launch list driver for ncu: runs a few ops eagerly, 3 times each, so `ncu --metrics gpu__time_duration.sum` lists
the device time of every kernel behind them.  python tools/prof_multi.py [N] [ops...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import flownet2_tf_b200 as fn2  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 32
which = sys.argv[2:] or ["ds_multi", "epe", "x3_fwd", "x3_bwd"]
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
lr = torch.nn.functional.leaky_relu
sizes = [(6, 8), (12, 16), (24, 32), (48, 64), (96, 128)]
flo = torch.randn(N, 384, 512, 2, device=dev, generator=g) * 8
preds = [torch.randn(N, h, w, 2, device=dev, generator=g) for h, w in sizes]
a = lr(torch.randn(N, 48, 64, 256, device=dev, generator=g), 0.1)
b = lr(torch.randn(N, 48, 64, 256, device=dev, generator=g), 0.1)
go = torch.randn(N, 48, 64, 441, device=dev, generator=g)
img = torch.rand(N, 384, 512, 3, device=dev, generator=g)
gi = torch.randn(N, 384, 512, 3, device=dev, generator=g)
fns = {
    "ds_multi": lambda: fn2.downsample_multi(flo, sizes),
    "epe": lambda: fn2.multiscale_epe(flo, preds, 0.05),
    "x3_fwd": lambda: fn2.correlation(a, b, 1, 20, 1, 2, 20, precision="tf32x3"),
    "x3_bwd": lambda: fn2.correlation_grad(go, a, b, 1, 20, 1, 2, 20, precision="tf32x3"),
    "tf32_fwd": lambda: fn2.correlation(a, b, 1, 20, 1, 2, 20, precision="tf32"),
    "tf32_bwd": lambda: fn2.correlation_grad(go, a, b, 1, 20, 1, 2, 20, precision="tf32"),
    "warp_fwd": lambda: fn2.flow_warp(img, flo),
    "warp_bwd": lambda: fn2.flow_warp_grad(img, flo, gi),
}
for name in which:
    for _ in range(3):
        fns[name]()
    torch.cuda.synchronize()
print("ok")
