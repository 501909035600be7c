"""Worker of tests/test_gpu_strips.py::test_nccl_strips_match_the_oracle (one process per GPU, launched with
torch.distributed.run): the row-strip correlation with real NCCL send/recv of the halos, every rank's strip of the
forward and of both gradients against the float64 CPU oracle of the WHOLE image (oracle/fn2_oracle.c, the restatement
of correlation_kernel.cu.cc:21-110 and correlation_grad_kernel.cu.cc:20-189).  Test infrastructure, not a product path.

    argv: precision ("fp32" | "tf32" | "tf32x3" | "default")   rows-per-strip   width   channels
The schedule (exchange overlapped with the interior rows, or exchange first) follows FN2_STRIPS_OVERLAP_MIN_ROWS."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import flownet2_tf_b200 as fn2  # noqa: E402,F401
from flownet2_tf_b200 import strips  # noqa: E402
from oracle import oracle as o  # noqa: E402
from conftest import assert_close  # noqa: E402

ARGS = (1, 20, 1, 2, 20)
TF32_TOL = 2.5e-3   # of the tensor's maximum (fn2ops.h, FN2_PREC_TF32)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    prec = None if sys.argv[1] == "default" else sys.argv[1]
    rows, W, C = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    H = rows * world
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    g = torch.Generator().manual_seed(4321)               # the same whole image on every rank
    lr = torch.nn.functional.leaky_relu
    a = lr(torch.randn(1, H, W, C, generator=g), 0.1)
    b = lr(torch.randn(1, H, W, C, generator=g), 0.1)
    go = torch.randn(1, H, W, 441, generator=g)
    want = o.c_correlation(a.numpy(), b.numpy(), *ARGS, f64=True)
    wga, wgb = o.c_correlation_grad(go.numpy(), a.numpy(), b.numpy(), *ARGS, f64=True)
    sl = slice(rank * rows, (rank + 1) * rows)
    a_s = a[:, sl].contiguous().to(dev).requires_grad_()
    b_s = b[:, sl].contiguous().to(dev).requires_grad_()
    out = strips.correlation_strips(a_s, b_s, *ARGS, precision=prec)
    out.backward(go[:, sl].contiguous().to(dev))
    torch.cuda.synchronize()
    worst = 0.0
    for name, got, ref in (("forward", out, want), ("backprops_a", a_s.grad, wga), ("backprops_b", b_s.grad, wgb)):
        got = got.detach().cpu().numpy().astype(np.float64)
        ref_s = ref[:, sl]
        rel = float(np.abs(got - ref_s).max() / max(np.abs(ref).max(), 1e-30))
        worst = max(worst, rel)
        if prec == "tf32":
            assert rel <= TF32_TOL, "%s rank %d: %.3e of max" % (name, rank, rel)
        else:   # fp32, split-tf32 and the default: the reference's own tolerance, elementwise
            assert_close(got, ref_s, rel=1e-5, abs_=1e-6, what="%s rank %d" % (name, rank))
    w = torch.tensor([worst], device=dev)
    dist.all_reduce(w, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"gpus": world, "precision": sys.argv[1], "rows_per_gpu": rows,
                          "overlap": bool(strips.overlap_used(1, rows, W, C, 20, 2, 0, world)),
                          "max_err_of_max_vs_f64_oracle": float(w.item())}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
