import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import flownet2_tf_b200 as fn2
from flownet2_tf_b200 import _lib
dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(0)
for shape in [(2, 12, 16, 256), (2, 12, 16, 64), (1, 14, 18, 32), (2, 24, 32, 256)]:
    a = torch.randn(*shape, generator=g).to(dev); b = torch.randn(*shape, generator=g).to(dev)
    go = torch.randn(shape[0], shape[1], shape[2], 441, generator=g).to(dev)
    for p in ("tf32", "tf32x3"):
        try:
            fn2.correlation(a, b, 1, 20, 1, 2, 20, precision=p)
            fn2.correlation_grad(go, a, b, 1, 20, 1, 2, 20, precision=p)
            torch.cuda.synchronize()
            print(shape, p, "ok")
        except Exception as e:
            print(shape, p, "FAIL", e, _lib.lib().fn2_last_cuda_error())
