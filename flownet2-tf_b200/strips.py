"""Row-strip partitioning of the correlation across GPUs (BASELINE config 5: 4K inputs).

The image height is cut into ``world`` horizontal strips, one per rank (rank 0 = top).  For FlowNetC's call
(kernel_size 1, stride_1 1, pad == max_displacement; ``src/flownet_c/flownet_c.py:40``) output row ``y`` needs
``a[y]`` and ``b[y - h .. y + h]`` with ``h = (max_displacement // stride_2) * stride_2`` (20 rows), so

  forward : every rank sends its top ``h`` rows of ``b`` to the rank above and its bottom ``h`` rows to the rank
            below (``torch.distributed`` P2P: NCCL send/recv over NVLink on GPUs, gloo in the CPU tests), then runs
            ``fn2_correlation_forward_halo`` on its own rows.  The image border uses no halo: rows beyond the tall
            ``b`` are the op's zero padding.
  backward: ``backprops_a`` needs the same halo of ``b`` (kept from the forward).  ``backprops_b`` is produced for
            the tall tensor; its halo rows are partial sums owned by the neighbours, so the mirror exchange sends
            them back and the owner adds them (2 x 9.83 MB per boundary at 4K instead of shipping 441-channel
            gradient halos).

The reference has no multi-GPU path at all (``src/net.py`` builds one graph on one device); this is the B200-side
extension BASELINE.json's north_star asks for.  Both steps are ``torch.autograd.Function``s, so
``correlation_strips`` is differentiable end to end.
"""
import torch
import torch.distributed as dist

from . import _lib
from .correlation import _check_pair, _prec, _workspace, output_shape


def halo_rows(max_displacement, stride_2):
    """Rows of ``b`` a strip needs from each neighbour: the largest vertical displacement, r * stride_2."""
    return (int(max_displacement) // int(stride_2)) * int(stride_2)


# --------------------------------------------------------------------------------------------- the local op
def _geom(a, k, md, s1, s2, pad):
    N, H, W, C = a.shape
    return (N, H, W, C, k, md, s1, s2, pad)


def correlation_halo_forward(a, b_ext, k, md, s1, s2, pad, halo_top, halo_bottom, precision=None):
    """out rows of one strip: ``a`` [N,H,W,C], ``b_ext`` [N,H+halo_top+halo_bottom,W,C] (fn2ops.h, *_halo)."""
    _lib.require_cuda_f32(a, "input_a")
    _lib.require_cuda_f32(b_ext, "input_b")
    N, H, W, C = a.shape
    prec = _prec(precision, _geom(a, k, md, s1, s2, pad))
    if tuple(b_ext.shape) != (N, H + halo_top + halo_bottom, W, C):
        raise ValueError("input_b must have shape %s, got %s" % ((N, H + halo_top + halo_bottom, W, C),
                                                                 tuple(b_ext.shape)))
    oh, ow, oc = output_shape(H, W, k, md, s1, s2, pad)
    out = torch.empty((N, oh, ow, oc), dtype=torch.float32, device=a.device)
    geom = _geom(a, k, md, s1, s2, pad)
    with torch.cuda.device(a.device):
        ws, nbytes, wsp = _workspace(a, geom, prec, 0)
        st = _lib.lib().fn2_correlation_forward_halo(a.data_ptr(), b_ext.data_ptr(), out.data_ptr(), *geom,
                                                     halo_top, halo_bottom, prec, wsp, nbytes,
                                                     _lib.current_stream_ptr(a.device))
    _lib.check(st, "Correlation(halo)")
    return out


def correlation_halo_backward(gradients, a, b_ext, k, md, s1, s2, pad, halo_top, halo_bottom, precision=None):
    """(backprops_a [N,H,W,C], backprops_b_ext [N,H+halo_top+halo_bottom,W,C]) of one strip."""
    for t, name in ((gradients, "gradients"), (a, "input_a"), (b_ext, "input_b")):
        _lib.require_cuda_f32(t, name)
    prec = _prec(precision, _geom(a, k, md, s1, s2, pad))
    ga = torch.empty_like(a)
    gb = torch.empty_like(b_ext)
    geom = _geom(a, k, md, s1, s2, pad)
    with torch.cuda.device(a.device):
        ws, nbytes, wsp = _workspace(a, geom, prec, 1)
        st = _lib.lib().fn2_correlation_backward_halo(gradients.data_ptr(), a.data_ptr(), b_ext.data_ptr(),
                                                      ga.data_ptr(), gb.data_ptr(), *geom, halo_top, halo_bottom,
                                                      prec, wsp, nbytes, _lib.current_stream_ptr(a.device))
    _lib.check(st, "CorrelationGrad(halo)")
    return ga, gb


class _CorrelationHalo(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b_ext, attrs, ops):
        ctx.save_for_backward(a, b_ext)
        ctx.attrs, ctx.ops = attrs, ops
        return ops[0](a, b_ext, *attrs)

    @staticmethod
    def backward(ctx, gradients):
        a, b_ext = ctx.saved_tensors
        ga, gb = ctx.ops[1](gradients.contiguous(), a, b_ext, *ctx.attrs)
        return ga, gb, None, None


_CUDA_OPS = (correlation_halo_forward, correlation_halo_backward)


# ------------------------------------------------------------------------------------------ the exchange step
def _neighbours(group):
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    to_global = (lambda r: dist.get_global_rank(group, r)) if group is not None else (lambda r: r)
    up = to_global(rank - 1) if rank > 0 else None
    down = to_global(rank + 1) if rank < world - 1 else None
    return up, down


def _exchange(send_up, send_down, up, down, group):
    """Send one tensor to each existing neighbour and receive one of the same shape from it."""
    recv_up = torch.empty_like(send_up) if up is not None else None
    recv_down = torch.empty_like(send_down) if down is not None else None
    ops = []
    if up is not None:
        ops += [dist.P2POp(dist.isend, send_up, up, group), dist.P2POp(dist.irecv, recv_up, up, group)]
    if down is not None:
        ops += [dist.P2POp(dist.isend, send_down, down, group), dist.P2POp(dist.irecv, recv_down, down, group)]
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return recv_up, recv_down


class _HaloExchange(torch.autograd.Function):
    """b_strip [N,H,W,C] -> b_ext [N, top + H + bottom, W, C] with the neighbours' boundary rows attached."""

    @staticmethod
    def forward(ctx, b, halo, group):
        up, down = _neighbours(group)
        if (up is not None or down is not None) and b.shape[1] < halo:
            raise ValueError("strip height %d is smaller than the halo (%d rows)" % (b.shape[1], halo))
        top_rows = b[:, :halo].contiguous()
        bottom_rows = b[:, b.shape[1] - halo:].contiguous()
        recv_up, recv_down = _exchange(top_rows, bottom_rows, up, down, group)
        ctx.halo, ctx.group, ctx.nb = halo, group, (up, down)
        parts = ([recv_up] if up is not None else []) + [b] + ([recv_down] if down is not None else [])
        return torch.cat(parts, dim=1) if len(parts) > 1 else b.clone()

    @staticmethod
    def backward(ctx, gb_ext):
        up, down = ctx.nb
        halo = ctx.halo
        top = halo if up is not None else 0
        bottom = halo if down is not None else 0
        H = gb_ext.shape[1] - top - bottom
        gb = gb_ext[:, top:top + H].clone()
        # partial sums for the neighbours' rows go back to their owners; theirs for our rows come in
        send_up = gb_ext[:, :top].contiguous() if up is not None else gb_ext[:, :0]
        send_down = gb_ext[:, top + H:].contiguous() if down is not None else gb_ext[:, :0]
        recv_up, recv_down = _exchange(send_up, send_down, up, down, ctx.group)
        if recv_up is not None:
            gb[:, :halo] += recv_up
        if recv_down is not None:
            gb[:, H - halo:] += recv_down
        return gb, None, None


def halo_sizes(group=None, max_displacement=20, stride_2=2):
    """(halo_top, halo_bottom) of this rank: 0 at the image border, r * stride_2 next to a neighbour."""
    up, down = _neighbours(group)
    h = halo_rows(max_displacement, stride_2)
    return (h if up is not None else 0), (h if down is not None else 0)


def correlation_strips(a_strip, b_strip, kernel_size, max_displacement, stride_1, stride_2, padding, *, group=None,
                       precision=None, local_ops=None):
    """Correlation of a row-partitioned pair: every rank passes ITS rows of input_a / input_b and gets its rows of
    the cost volume.  Same positional arguments as ``correlation``.  ``local_ops`` (forward, backward callables with
    the signature of ``correlation_halo_forward/backward``) exists for the CPU ``gloo`` tests of the exchange logic."""
    args = tuple(int(v) for v in (kernel_size, max_displacement, stride_1, stride_2, padding))
    k, md, s1, s2, pad = args
    if k != 1 or s1 != 1 or pad != md:
        raise ValueError("strip partitioning needs kernel_size 1, stride_1 1 and padding == max_displacement")
    if a_strip.shape != b_strip.shape:
        raise ValueError("input_a and input_b strips must have the same shape")
    h = halo_rows(md, s2)
    top, bottom = halo_sizes(group, md, s2)
    b_ext = _HaloExchange.apply(b_strip, h, group)
    if local_ops is None:
        _check_pair(a_strip, b_strip)
        local_ops = _CUDA_OPS
    return _CorrelationHalo.apply(a_strip, b_ext, args + (top, bottom, precision), local_ops)
