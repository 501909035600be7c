"""Host logic of the row-strip partitioning (flownet2_tf_b200.strips) on CPU: world_size 2 and 3 over gloo, with the
oracle standing in for the CUDA kernel (``local_ops``).  Checks that halo exchange + local op + mirror exchange of the
b-gradient halos reproduce the whole-image op and its gradients."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import corr_inputs, grad_like, oracle_halo_backward, oracle_halo_forward

ARGS = (1, 20, 1, 2, 20)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _bounds(H, world):
    base, rem = divmod(H, world)
    edges = [0]
    for r in range(world):
        edges.append(edges[-1] + base + (1 if r < rem else 0))
    return edges


def _worker(rank, world, port, H, W, C, outdir, overlap):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import flownet2_tf_b200.strips as strips
        strips.OVERLAP_MIN_ROWS = 0 if overlap else 10 ** 9    # force either schedule, whatever the strip height
        a, b = corr_inputs(1, H, W, C)
        e = _bounds(H, world)
        a_s = a[:, e[rank]:e[rank + 1]].clone().requires_grad_()
        b_s = b[:, e[rank]:e[rank + 1]].clone().requires_grad_()
        out = strips.correlation_strips(a_s, b_s, *ARGS, local_ops=(oracle_halo_forward, oracle_halo_backward))
        g = grad_like((1, H, W, 441))[:, e[rank]:e[rank + 1]].contiguous()
        out.backward(g)
        np.savez(os.path.join(outdir, "r%d.npz" % rank), out=out.detach().numpy(), ga=a_s.grad.numpy(),
                 gb=b_s.grad.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,H,overlap", [
    (2, 44, True), (3, 66, True), (2, 41, True),      # strips shorter than 2 halos: exchange, then one call, always
    (2, 100, True), (3, 150, True),                   # interior rows computed under the exchange
    (2, 100, False)])                                 # the same strips, exchange first and one call
def test_strips_reproduce_whole_image_op(tmp_path, world, H, overlap):
    from oracle import oracle as o
    W, C = 10, 4
    port = _free_port()
    mp.spawn(_worker, args=(world, port, H, W, C, str(tmp_path), overlap), nprocs=world, join=True)
    a, b = corr_inputs(1, H, W, C)
    g = grad_like((1, H, W, 441))
    want = o.c_correlation(a.numpy(), b.numpy(), *ARGS, f64=True)
    wga, wgb = o.c_correlation_grad(g.numpy(), a.numpy(), b.numpy(), *ARGS, f64=True)
    parts = [np.load(os.path.join(str(tmp_path), "r%d.npz" % r)) for r in range(world)]
    np.testing.assert_allclose(np.concatenate([p["out"] for p in parts], 1), want, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(np.concatenate([p["ga"] for p in parts], 1), wga, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(np.concatenate([p["gb"] for p in parts], 1), wgb, rtol=1e-5, atol=2e-6)


def test_strip_shorter_than_halo_is_rejected():
    import flownet2_tf_b200.strips as strips
    assert strips.halo_rows(20, 2) == 20 and strips.halo_rows(21, 2) == 20 and strips.halo_rows(4, 1) == 4
    with pytest.raises(ValueError, match="kernel_size 1"):
        strips.correlation_strips(torch.zeros(1, 4, 4, 4), torch.zeros(1, 4, 4, 4), 3, 20, 1, 2, 20)
