"""flow_warp forward / backward device time on synthetic flows (N(0, sigma) px) and on the reference's real sample flow
(tests/golden/sample0.npz, rolled per image).  python tools/time_warp.py [N]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import flownet2_tf_b200 as fn2  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
ROT = 3


def timed(fn):
    for i in range(ROT):
        fn(i)
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        with torch.cuda.graph(gr, stream=s):
            for i in range(ROT):
                fn(i)
    for _ in range(3):
        gr.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gr.replay(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / ROT)
    return sorted(ts)[2]


z = np.load(os.path.join(ROOT, "tests", "golden", "sample0.npz"))
img1 = torch.from_numpy(z["image"].astype(np.float32)).to(dev)
flo1 = torch.from_numpy(z["flow"]).to(dev)
cases = {}
for sigma in (8.0, 2.0):
    cases["sigma%g" % sigma] = ([torch.rand(N, 384, 512, 3, device=dev, generator=g) for _ in range(ROT)],
                                [torch.randn(N, 384, 512, 2, device=dev, generator=g) * sigma for _ in range(ROT)])
cases["sample_flow"] = ([torch.stack([img1.roll(7 * (i * N + j), 0) for j in range(N)]).contiguous() for i in range(ROT)],
                        [torch.stack([flo1.roll(7 * (i * N + j), 0) for j in range(N)]).contiguous() for i in range(ROT)])
GI = [torch.randn(N, 384, 512, 3, device=dev, generator=g) for _ in range(ROT)]
for name, (IMG, FLO) in cases.items():
    f = timed(lambda i: fn2.flow_warp(IMG[i], FLO[i]))
    b = timed(lambda i: fn2.flow_warp_grad(IMG[i], FLO[i], GI[i]))
    print(json.dumps({"flow": name, "N": N, "warp_fwd_ms": round(f, 4), "warp_bwd_ms": round(b, 4)}), flush=True)
