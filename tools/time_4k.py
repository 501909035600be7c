"""Device time of the correlation at the 4K conv3 shape (cfg5: [1,272,480,256]) -- whole image on one GPU and one of G
row strips with its halos (what each rank of the strip partition runs).  python tools/time_4k.py [precision]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import flownet2_tf_b200 as fn2  # noqa: E402
from flownet2_tf_b200 import strips  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "tf32"
ARGS = (1, 20, 1, 2, 20)
H, W, C, D = 272, 480, 256, 441
dev = torch.device("cuda", 0)
HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
g = torch.Generator(device=dev).manual_seed(3)
lr = torch.nn.functional.leaky_relu


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def report(name, ms, nbytes):
    gbs = nbytes / ms / 1e6
    print(json.dumps({"op": name, "precision": prec, "ms": round(ms, 4), "alg_mb": round(nbytes / 1e6, 1),
                      "gb_s": round(gbs, 1), "hbm_frac": round(gbs / HBM, 4)}), flush=True)


for G in (1, 2, 4, 8):
    Hs = H // G
    top = bottom = 20 if G > 1 else 0     # an interior strip: both halos
    a = lr(torch.randn(1, Hs, W, C, device=dev, generator=g), 0.1)
    b = lr(torch.randn(1, Hs + top + bottom, W, C, device=dev, generator=g), 0.1)
    go = torch.randn(1, Hs, W, D, device=dev, generator=g)
    act, vol = Hs * W * C * 4, Hs * W * D * 4
    ext = (Hs + top + bottom) * W * C * 4
    ms = timed(lambda: strips.correlation_halo_forward(a, b, *ARGS, top, bottom, prec))
    report("corr_fwd_4k_strip_1of%d" % G, ms, act + ext + vol)
    ms = timed(lambda: strips.correlation_halo_backward(go, a, b, *ARGS, top, bottom, prec))
    report("corr_bwd_4k_strip_1of%d" % G, ms, vol + 2 * act + 2 * ext)
