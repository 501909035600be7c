"""Tiny driver for ncu: python tools/prof_op.py <op> [N]; op in corr_fwd corr_bwd warp_fwd warp_bwd ds<oh>x<ow>."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import flownet2_tf_b200 as fn2  # noqa: E402

op = sys.argv[1]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 32
prec = sys.argv[3] if len(sys.argv) > 3 else "tf32"
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
lr = torch.nn.functional.leaky_relu
if op.startswith("corr"):
    a = lr(torch.randn(N, 48, 64, 256, device=dev, generator=g), 0.1)
    b = lr(torch.randn(N, 48, 64, 256, device=dev, generator=g), 0.1)
    go = torch.randn(N, 48, 64, 441, device=dev, generator=g)
    fn = (lambda: fn2.correlation(a, b, 1, 20, 1, 2, 20, precision=prec)) if op == "corr_fwd" else \
         (lambda: fn2.correlation_grad(go, a, b, 1, 20, 1, 2, 20, precision=prec))
else:
    img = torch.rand(N, 384, 512, 3, device=dev, generator=g)
    flo = torch.randn(N, 384, 512, 2, device=dev, generator=g) * 8
    gi = torch.randn(N, 384, 512, 3, device=dev, generator=g)
    if op == "warp_fwd":
        fn = lambda: fn2.flow_warp(img, flo)
    elif op == "warp_bwd":
        fn = lambda: fn2.flow_warp_grad(img, flo, gi)
    elif op == "ds_multi":
        fn = lambda: fn2.downsample_multi(flo, [(6, 8), (12, 16), (24, 32), (48, 64), (96, 128)])
    else:
        oh, ow = [int(v) for v in op[2:].split("x")]
        fn = lambda: fn2.downsample(flo, [oh, ow])
for _ in range(4):
    fn()
torch.cuda.synchronize()
print("ok")
