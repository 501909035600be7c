"""Developer check for the tensor-core correlation: error statistics against the f64 oracle and kernel timings.
Run on a GPU box:  python tools/check_tc.py [precision ...]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import flownet2_tf_b200 as fn2  # noqa: E402
from oracle import oracle as o  # noqa: E402
from conftest import corr_inputs, grad_like  # noqa: E402

ARGS = (1, 20, 1, 2, 20)


def stats(name, got, want):
    err = np.abs(got.astype(np.float64) - want)
    rel = err / (np.abs(want) + 1e-6)
    tol = err / (1e-6 + 1e-5 * np.abs(want))
    print("  %-12s max_abs %.3e  max_rel %.3e  mean_abs %.3e  worst/tol(1e-5,1e-6) %.2f  |want|max %.3f"
          % (name, err.max(), rel.max(), err.mean(), tol.max(), np.abs(want).max()), flush=True)
    return err


def check(prec, shape, do_bwd=True):
    N, H, W, C = shape
    a, b = corr_inputs(N, H, W, C)
    print("precision=%s shape=%s" % (prec, shape), flush=True)
    try:
        out = fn2.correlation(a.cuda(), b.cuda(), *ARGS, precision=prec)
        torch.cuda.synchronize()
    except Exception as e:
        print("  forward FAILED:", repr(e)[:300])
        return
    want = o.c_correlation(a.numpy(), b.numpy(), *ARGS, f64=True)
    err = stats("forward", out.cpu().numpy(), want)
    if err.max() > 1e-2:
        bad = np.argwhere(err > 1e-2)
        print("  bad count", len(bad), "first", bad[:5].tolist(), "last", bad[-3:].tolist())
        e4 = err.reshape(N, H, W, 21, 21)
        print("  err by y%2,x%2:", [[float(e4[:, i::2, j::2].max()) for j in range(2)] for i in range(2)])
        print("  err by y:", np.round(e4.max(axis=(0, 2, 3, 4)), 3).tolist())
        print("  err by p:", np.round(e4.max(axis=(0, 1, 2, 4)), 3).tolist())
        print("  err by o:", np.round(e4.max(axis=(0, 1, 2, 3)), 3).tolist())
    if do_bwd:
        g = grad_like(out.shape)
        try:
            ga, gb = fn2.correlation_grad(g.cuda(), a.cuda(), b.cuda(), *ARGS, precision=prec)
            torch.cuda.synchronize()
        except Exception as e:
            print("  backward FAILED:", repr(e)[:300])
            return
        wga, wgb = o.c_correlation_grad(g.numpy(), a.numpy(), b.numpy(), *ARGS, f64=True)
        for name, got, want_g in (("backprops_a", ga.cpu().numpy(), wga), ("backprops_b", gb.cpu().numpy(), wgb)):
            e = stats(name, got, want_g)
            if e.max() > 2e-3 * np.abs(want_g).max():
                print("   err by y%2,x%2:", [[float(e[:, i::2, j::2].max()) for j in range(2)] for i in range(2)])
                print("   err by y:", np.round(e.max(axis=(0, 2, 3)) / np.abs(want_g).max(), 2).tolist())
                print("   err by x:", np.round(e.max(axis=(0, 1, 3)) / np.abs(want_g).max(), 2).tolist())
                print("   err by c/8:", np.round(e.reshape(N, H, W, -1, 8).max(axis=(0, 1, 2, 4)) / np.abs(want_g).max(), 2).tolist())
                print("   got[0,0,0,:8]", got[0, 0, 0, :8].tolist(), "want", want_g[0, 0, 0, :8].tolist())
                print("   got[0,20,30,:8]", got[0, 20 % H, 30 % W, :8].tolist(), "want", want_g[0, 20 % H, 30 % W, :8].tolist())


def timeit(prec, N=32, reps=20):
    a, b = corr_inputs(N, 48, 64, 256)
    a, b = a.cuda(), b.cuda()
    g = grad_like((N, 48, 64, 441)).cuda()
    try:
        for _ in range(3):
            out = fn2.correlation(a, b, *ARGS, precision=prec)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            out = fn2.correlation(a, b, *ARGS, precision=prec)
        e1.record()
        torch.cuda.synchronize()
        f = e0.elapsed_time(e1) / reps
        for _ in range(2):
            fn2.correlation_grad(g, a, b, *ARGS, precision=prec)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            fn2.correlation_grad(g, a, b, *ARGS, precision=prec)
        e1.record()
        torch.cuda.synchronize()
        bw = e0.elapsed_time(e1) / 5
        fb, bb = N * 11.710e6, N * 18.002e6
        print("TIME %s N=%d: fwd %.3f ms (%.0f GB/s, %.1f%% of 6459.6)  bwd %.3f ms (%.0f GB/s, %.1f%%)"
              % (prec, N, f, fb / f / 1e6, fb / f / 1e6 / 64.596, bw, bb / bw / 1e6, bb / bw / 1e6 / 64.596), flush=True)
    except Exception as e:
        print("TIME %s FAILED: %r" % (prec, repr(e)[:300]))


if __name__ == "__main__":
    precs = sys.argv[1:] or ["tf32"]
    for prec in precs:
        check(prec, (1, 48, 64, 256))
        check(prec, (2, 13, 18, 64))
        check(prec, (3, 47, 61, 96))
        timeit(prec)
    timeit("fp32", reps=3)
