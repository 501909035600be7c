"""Device time of every op of the hot path, free of Python launch overhead: each op is captured in a CUDA graph that
runs it over ROT distinct input sets (so consecutive launches never find their inputs in L2) and the graph replay is
timed with CUDA events.  python tools/time_ops.py [N] [precision]  -> one JSON line per op."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import flownet2_tf_b200 as fn2  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 32
prec = sys.argv[2] if len(sys.argv) > 2 else "tf32"
only = sys.argv[3].split(",") if len(sys.argv) > 3 else None
ROT = 4
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
lr = torch.nn.functional.leaky_relu
HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def rnd(*s):
    return torch.randn(*s, device=dev, generator=g)


def timed(name, fn, nbytes, reps=5):
    for i in range(ROT):
        fn(i)
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        fn(0)
        torch.cuda.synchronize()
        with torch.cuda.graph(gr, stream=s):
            for i in range(ROT):
                fn(i)
    for _ in range(3):
        gr.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        gr.replay()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / ROT)
    ts.sort()
    ms = ts[len(ts) // 2]
    gbs = nbytes / (ms * 1e-3) / 1e9
    print(json.dumps({"op": name, "N": N, "ms": round(ms, 4), "min_ms": round(ts[0], 4), "alg_mb": round(nbytes / 1e6, 2),
                      "gb_s": round(gbs, 1), "hbm_frac": round(gbs / HBM, 4)}), flush=True)


def want(n):
    return only is None or any(n.startswith(o) or o.startswith(n) for o in only)


f = 4
if want("corr"):
    A = [lr(rnd(N, 48, 64, 256), 0.1) for _ in range(ROT)]
    B = [lr(rnd(N, 48, 64, 256), 0.1) for _ in range(ROT)]
    G = [rnd(N, 48, 64, 441) for _ in range(ROT)]
    act, vol = N * 48 * 64 * 256 * f, N * 48 * 64 * 441 * f
    if want("corr_fwd"):
        timed("corr_fwd", lambda i: fn2.correlation(A[i], B[i], 1, 20, 1, 2, 20, precision=prec), 2 * act + vol)
    if want("corr_bwd"):
        timed("corr_bwd", lambda i: fn2.correlation_grad(G[i], A[i], B[i], 1, 20, 1, 2, 20, precision=prec), vol + 4 * act)
    del A, B, G
if want("warp") or want("down"):
    IMG = [torch.rand(N, 384, 512, 3, device=dev, generator=g) for _ in range(ROT)]
    FLO = [rnd(N, 384, 512, 2) * 8 for _ in range(ROT)]
    GI = [rnd(N, 384, 512, 3) for _ in range(ROT)]
    img, flo = N * 384 * 512 * 3 * f, N * 384 * 512 * 2 * f
    if want("warp_fwd"):
        timed("warp_fwd", lambda i: fn2.flow_warp(IMG[i], FLO[i]), 2 * img + flo)
    if want("warp_bwd"):
        timed("warp_bwd", lambda i: fn2.flow_warp_grad(IMG[i], FLO[i], GI[i]), 4 * img + 2 * flo)
    if want("down"):
        for s in [(6, 8), (12, 16), (24, 32), (48, 64), (96, 128)]:
            timed("downsample_%dx%d" % s, lambda i, s=s: fn2.downsample(FLO[i], list(s)), flo + N * s[0] * s[1] * 2 * f)
        sizes = [(6, 8), (12, 16), (24, 32), (48, 64), (96, 128)]
        timed("downsample_multi", lambda i: fn2.downsample_multi(FLO[i], sizes),
              flo + sum(N * h * w * 2 * f for h, w in sizes))
        PR = [rnd(N, h, w, 2) for h, w in sizes]
        timed("multiscale_epe", lambda i: fn2.multiscale_epe(FLO[i], PR, 0.05),
              flo + sum(N * h * w * 2 * f for h, w in sizes))
