"""Generates tests/golden/*.npz: outputs of the REFERENCE'S OWN CUDA kernels (oracle/_ref/libfn2ref.so, compiled
unmodified from /root/reference/src/ops by oracle/Makefile) on small seeded inputs.  Run on a GPU box:

    python tests/golden/make_golden.py gpurun_out/golden      # then copy the .npz files into tests/golden/

The inputs are stored next to the outputs, so the CPU tests (tests/test_golden.py) need neither a GPU nor the
reference checkout.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref  # noqa: E402
from conftest import corr_inputs, grad_like, warp_inputs  # noqa: E402

CORR = {
    "corr_flownetc": (1, 14, 18, 32, 1, 20, 1, 2, 20),
    # H*W % 16 == 0 below: keeps the reference's PadData out of its buggy early-return branch (pad.cu.cc:31-33)
    "corr_k3_s2": (1, 8, 12, 5, 3, 3, 2, 2, 3),
    "corr_padgt": (1, 12, 12, 6, 3, 2, 1, 1, 4),
    "corr_pwc": (2, 8, 12, 16, 1, 4, 1, 1, 4),
}


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    dev = torch.device("cuda", 0)
    for name, (N, H, W, C, k, md, s1, s2, pad) in CORR.items():
        a, b = corr_inputs(N, H, W, C)
        out = ref.correlation(a.to(dev), b.to(dev), k, md, s1, s2, pad)
        g = grad_like(out.shape)
        ga, gb = ref.correlation_grad(g.to(dev), a.to(dev), b.to(dev), k, md, s1, s2, pad)
        torch.cuda.synchronize()
        np.savez_compressed(os.path.join(outdir, name + ".npz"), args=np.array([k, md, s1, s2, pad]), a=a.numpy(),
                            b=b.numpy(), g=g.numpy(), out=out.cpu().numpy(), ga=ga.cpu().numpy(), gb=gb.cpu().numpy())
    image, flow = warp_inputs(1, 24, 30, 3)
    g = grad_like(image.shape)
    w = ref.flow_warp(image.to(dev), flow.to(dev))
    gi, gf = ref.flow_warp_grad(image.to(dev), flow.to(dev), g.to(dev))
    torch.cuda.synchronize()
    np.savez_compressed(os.path.join(outdir, "warp.npz"), image=image.numpy(), flow=flow.numpy(), g=g.numpy(),
                        out=w.cpu().numpy(), gi=gi.cpu().numpy(), gf=gf.cpu().numpy())
    gen = torch.Generator().manual_seed(1237)
    x = torch.randn(1, 48, 64, 2, generator=gen) * 5
    x[0, 10:20, 10:30] = float("nan")
    outs = {"x": x.numpy()}
    for size in [(6, 8), (12, 16), (24, 32), (5, 7)]:
        outs["out_%d_%d" % size] = ref.downsample(x.to(dev), size).cpu().numpy()
    torch.cuda.synchronize()
    np.savez_compressed(os.path.join(outdir, "downsample.npz"), **outs)
    print("golden vectors written to", outdir)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
