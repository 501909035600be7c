"""Row-strip partitioning of the correlation across GPUs (BASELINE config 5: 4K inputs).

The image height is cut into ``world`` horizontal strips, one per rank (rank 0 = top).  For FlowNetC's call
(kernel_size 1, stride_1 1, pad == max_displacement; ``src/flownet_c/flownet_c.py:40``) output row ``y`` needs
``a[y]`` and ``b[y - h .. y + h]`` with ``h = (max_displacement // stride_2) * stride_2`` (20 rows), so

  forward : every rank sends its top ``h`` rows of ``b`` to the rank above and its bottom ``h`` rows to the rank
            below (``torch.distributed`` P2P: NCCL send/recv over NVLink on GPUs, gloo in the CPU tests); the halos
            land directly in a tall copy of ``b`` (no concatenation).  For strips of at least ``OVERLAP_MIN_ROWS``
            (100) rows and a batch of one, the rows that need no halo -- all but the first and last ``h`` of the
            strip -- are computed WHILE the exchange is in flight and the two boundary bands follow once the halos
            have arrived (``fn2_correlation_forward_halo`` on row slices).  Shorter strips (the 68- and 34-row strips
            of a 4K pair on 4 and 8 GPUs) wait for the exchange and run ONE kernel over the whole strip, which
            measured faster there; batches above one do the same and also stage the halos through contiguous
            buffers.  ``overlap_used`` tells which schedule a shape gets.  The image border uses no halo: rows
            beyond the tall ``b`` are the op's zero padding.
  backward: ``backprops_a`` needs the same halo of ``b`` (kept from the forward).  With the overlapped schedule the
            boundary bands are computed FIRST: their ``backprops_b`` covers the neighbours' halo rows, partial sums the owners must add, so the
            mirror exchange starts as soon as the bands are done and the interior rows are computed under it
            (2 x 9.83 MB per boundary at 4K instead of shipping 441-channel gradient halos).

With NCCL every step only enqueues work (sends / receives on NCCL's stream, kernels on the current stream, ordered by
events), so the host never blocks and the whole step can be captured in a CUDA graph.  The reference has no multi-GPU
path at all (``src/net.py`` builds one graph on one device); this is the B200-side extension BASELINE.json's
north_star asks for.  ``correlation_strips`` is a ``torch.autograd.Function``: differentiable end to end.
"""
import os

import torch
import torch.distributed as dist

from . import _lib
from .correlation import _check_pair, _prec, _workspace, output_shape


#: Strips shorter than this run the exchange first and ONE kernel over the whole strip instead of overlapping the
#: exchange with the interior rows (three kernels per pass).  Overlap only pays when the interior is a launch of its
#: own worth: at the 4K conv3 shape (strips of 136 / 68 / 34 rows) the split costs more than the 0.05 - 0.14 ms exchange
#: it hides: 4 strips of 68 rows 0.61 ms plain against 0.67 overlapped (split-tf32), 2 strips of 136 rows 0.88 against 0.87
#: (bench.py --config strips4k).
OVERLAP_MIN_ROWS = int(os.environ.get("FN2_STRIPS_OVERLAP_MIN_ROWS", "100"))


def _overlap_decision(N, H, W, C, h, s2, top, bottom):
    """True when a strip of this shape overlaps the exchange with its interior rows: a batch of one (row slices stay
    contiguous), 2h own rows for a boundary band, an interior worth a launch, row slices that stay 16-byte aligned, and
    at least OVERLAP_MIN_ROWS rows (shorter strips measured faster with one kernel after the exchange)."""
    i0, i1 = (h if top else 0), (H - h if bottom else H)
    return bool(N == 1 and (top or bottom) and H >= 2 * h and i1 - i0 >= 2 * s2 and (W * C) % 4 == 0
                and H >= OVERLAP_MIN_ROWS)


def overlap_used(N, H, W, C, max_displacement, stride_2, rank, world):
    """Whether ``correlation_strips`` overlaps the halo exchange with the interior rows for a [N,H,W,C] strip held by
    ``rank`` of ``world`` (False: exchange first, then one kernel over the whole strip).  bench.py reports it."""
    h = halo_rows(max_displacement, stride_2)
    return _overlap_decision(N, H, W, C, h, stride_2, h if rank > 0 else 0, h if rank < world - 1 else 0)


def halo_rows(max_displacement, stride_2):
    """Rows of ``b`` a strip needs from each neighbour: the largest vertical displacement, r * stride_2."""
    return (int(max_displacement) // int(stride_2)) * int(stride_2)


# --------------------------------------------------------------------------------------------- the local op
def _geom(a, k, md, s1, s2, pad):
    N, H, W, C = a.shape
    return (N, H, W, C, k, md, s1, s2, pad)


def correlation_halo_forward(a, b_ext, k, md, s1, s2, pad, halo_top, halo_bottom, precision=None, out=None):
    """out rows of one strip: ``a`` [N,H,W,C], ``b_ext`` [N,H+halo_top+halo_bottom,W,C] (fn2ops.h, *_halo).
    ``out`` (optional) is a contiguous [N,H,W,D] tensor to write into (e.g. a row slice of a larger result)."""
    _lib.require_cuda_f32(a, "input_a")
    _lib.require_cuda_f32(b_ext, "input_b")
    N, H, W, C = a.shape
    prec = precision if isinstance(precision, int) else _prec(precision, _geom(a, k, md, s1, s2, pad))
    if tuple(b_ext.shape) != (N, H + halo_top + halo_bottom, W, C):
        raise ValueError("input_b must have shape %s, got %s" % ((N, H + halo_top + halo_bottom, W, C),
                                                                 tuple(b_ext.shape)))
    oh, ow, oc = output_shape(H, W, k, md, s1, s2, pad)
    if out is None:
        out = torch.empty((N, oh, ow, oc), dtype=torch.float32, device=a.device)
    else:
        _lib.require_cuda_f32(out, "output")
        if tuple(out.shape) != (N, oh, ow, oc):
            raise ValueError("output must have shape %s" % ((N, oh, ow, oc),))
    geom = _geom(a, k, md, s1, s2, pad)
    with torch.cuda.device(a.device):
        ws, nbytes, wsp = _workspace(a, geom, prec, 0)
        st = _lib.lib().fn2_correlation_forward_halo(a.data_ptr(), b_ext.data_ptr(), out.data_ptr(), *geom,
                                                     halo_top, halo_bottom, prec, wsp, nbytes,
                                                     _lib.current_stream_ptr(a.device))
    _lib.check(st, "Correlation(halo)")
    return out


def correlation_halo_backward(gradients, a, b_ext, k, md, s1, s2, pad, halo_top, halo_bottom, precision=None,
                              ga=None, gb=None):
    """(backprops_a [N,H,W,C], backprops_b_ext [N,H+halo_top+halo_bottom,W,C]) of one strip; ``ga`` / ``gb`` may be
    given as contiguous tensors (row slices of larger ones) to write into."""
    for t, name in ((gradients, "gradients"), (a, "input_a"), (b_ext, "input_b")):
        _lib.require_cuda_f32(t, name)
    prec = precision if isinstance(precision, int) else _prec(precision, _geom(a, k, md, s1, s2, pad))
    ga = torch.empty_like(a) if ga is None else _lib.require_cuda_f32(ga, "backprops_a")
    gb = torch.empty_like(b_ext) if gb is None else _lib.require_cuda_f32(gb, "backprops_b")
    if ga.shape != a.shape or gb.shape != b_ext.shape:
        raise ValueError("backprops_a / backprops_b must have the shapes of input_a / input_b")
    geom = _geom(a, k, md, s1, s2, pad)
    with torch.cuda.device(a.device):
        ws, nbytes, wsp = _workspace(a, geom, prec, 1)
        st = _lib.lib().fn2_correlation_backward_halo(gradients.data_ptr(), a.data_ptr(), b_ext.data_ptr(),
                                                      ga.data_ptr(), gb.data_ptr(), *geom, halo_top, halo_bottom,
                                                      prec, wsp, nbytes, _lib.current_stream_ptr(a.device))
    _lib.check(st, "CorrelationGrad(halo)")
    return ga, gb


def _cpu_adapter(ops):
    """Gives the two callables of the CPU gloo tests (plain halo signature, fresh outputs) the out= / ga= / gb=
    keywords of the CUDA ops."""
    fwd0, bwd0 = ops

    def fwd(a, b_ext, *args, out=None):
        r = fwd0(a, b_ext, *args)
        if out is None:
            return r
        out.copy_(r)
        return out

    def bwd(g, a, b_ext, *args, ga=None, gb=None):
        ra, rb = bwd0(g, a, b_ext, *args)
        if ga is not None:
            ga.copy_(ra)
            ra = ga
        if gb is not None:
            gb.copy_(rb)
            rb = gb
        return ra, rb
    return fwd, bwd


_CUDA_OPS = (correlation_halo_forward, correlation_halo_backward)


# ------------------------------------------------------------------------------------------ the exchange step
def _neighbours(group):
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    to_global = (lambda r: dist.get_global_rank(group, r)) if group is not None else (lambda r: r)
    up = to_global(rank - 1) if rank > 0 else None
    down = to_global(rank + 1) if rank < world - 1 else None
    return up, down


def _start_exchange(send_up, recv_up, send_down, recv_down, up, down, group):
    """Post the sends and receives with the existing neighbours and return the requests WITHOUT waiting.  With NCCL
    the transfers run on NCCL's own stream after the work already enqueued on the current stream; ``_finish`` makes
    the current stream wait for them (no host block), so kernels launched in between overlap the exchange."""
    ops = []
    if up is not None:
        ops += [dist.P2POp(dist.isend, send_up, up, group), dist.P2POp(dist.irecv, recv_up, up, group)]
    if down is not None:
        ops += [dist.P2POp(dist.isend, send_down, down, group), dist.P2POp(dist.irecv, recv_down, down, group)]
    return dist.batch_isend_irecv(ops) if ops else []


def _finish(reqs):
    for r in reqs:
        r.wait()


def halo_sizes(group=None, max_displacement=20, stride_2=2):
    """(halo_top, halo_bottom) of this rank: 0 at the image border, r * stride_2 next to a neighbour."""
    up, down = _neighbours(group)
    h = halo_rows(max_displacement, stride_2)
    return (h if up is not None else 0), (h if down is not None else 0)


def alloc_b_ext(N, H, W, C, device, group=None, max_displacement=20, stride_2=2):
    """A tall buffer [N, top + H + bottom, W, C] for this rank's ``input_b`` and the view of its own H rows.  A
    producer that writes ``input_b`` straight into the view (and passes ``b_ext=`` to ``correlation_strips``) saves
    the copy of the strip into the tall tensor."""
    top, bottom = halo_sizes(group, max_displacement, stride_2)
    b_ext = torch.empty((N, top + H + bottom, W, C), dtype=torch.float32, device=device)
    return b_ext, b_ext[:, top:top + H]


def _rows(t, y0, y1):
    """Row slice [y0, y1) of an NHWC tensor as a CONTIGUOUS tensor: a view when N == 1, a copy otherwise."""
    s = t[:, y0:y1]
    return s if s.is_contiguous() else s.contiguous()


class _CorrelationStrips(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, b_ext, attrs, group, ops):
        k, md, s1, s2, pad, precision = attrs
        fwd = ops[0]
        up, down = _neighbours(group)
        h = halo_rows(md, s2)
        top, bottom = (h if up is not None else 0), (h if down is not None else 0)
        N, H, W, C = a.shape
        if (up is not None or down is not None) and H < h:
            raise ValueError("strip height %d is smaller than the halo (%d rows)" % (H, h))
        if b_ext is None:
            b_ext = torch.empty((N, top + H + bottom, W, C), dtype=b.dtype, device=b.device)
            b_ext[:, top:top + H].copy_(b)
        own = b_ext[:, top:top + H]
        # boundary rows go out; the neighbours' rows come straight into the tall tensor (N == 1: the row slices are
        # contiguous, no staging)
        direct = N == 1
        send_up = _rows(own, 0, h) if up is not None else None
        send_down = _rows(own, H - h, H) if down is not None else None
        recv_up = (b_ext[:, :top] if direct else torch.empty((N, top, W, C), dtype=b.dtype, device=b.device)) \
            if up is not None else None
        recv_down = (b_ext[:, top + H:] if direct else torch.empty((N, bottom, W, C), dtype=b.dtype, device=b.device)) \
            if down is not None else None
        reqs = _start_exchange(send_up, recv_up, send_down, recv_down, up, down, group)
        oh, ow, oc = output_shape(H, W, k, md, s1, s2, pad)
        out = torch.empty((N, oh, ow, oc), dtype=a.dtype, device=a.device)
        # rows [i0, i1) need no halo: computed while the exchange is in flight
        i0, i1 = (h if top else 0), (H - h if bottom else H)
        overlap = bool(reqs) and _overlap_decision(N, H, W, C, h, s2, top, bottom)
        if overlap:
            fwd(_rows(a, i0, i1), _rows(own, i0 - (h if top else 0), i1 + (h if bottom else 0)), k, md, s1, s2, pad,
                h if top else 0, h if bottom else 0, precision, out=_rows(out, i0, i1))
        _finish(reqs)
        if not direct:
            if up is not None:
                b_ext[:, :top].copy_(recv_up)
            if down is not None:
                b_ext[:, top + H:].copy_(recv_down)
        if overlap:
            if top:      # rows [0, h): b_ext rows [0, top + 2h)
                fwd(_rows(a, 0, h), _rows(b_ext, 0, top + 2 * h), k, md, s1, s2, pad, top, h, precision,
                    out=_rows(out, 0, h))
            if bottom:   # rows [H - h, H): b_ext rows [top + H - 2h, end)
                fwd(_rows(a, H - h, H), _rows(b_ext, top + H - 2 * h, top + H + bottom), k, md, s1, s2, pad, h, bottom,
                    precision, out=_rows(out, H - h, H))
        else:
            fwd(a, b_ext, k, md, s1, s2, pad, top, bottom, precision, out=out)
        ctx.save_for_backward(a, b_ext)
        ctx.meta = (attrs, group, ops, up, down, h, top, bottom, overlap)
        return out

    @staticmethod
    def backward(ctx, gradients):
        a, b_ext = ctx.saved_tensors
        (k, md, s1, s2, pad, precision), group, ops, up, down, h, top, bottom, overlap = ctx.meta
        bwd = ops[1]
        N, H, W, C = a.shape
        gradients = gradients.contiguous()
        ga = torch.empty_like(a)
        gb = torch.empty((N, H, W, C), dtype=a.dtype, device=a.device)      # this rank's own rows of backprops_b
        recv_up = torch.empty((N, h, W, C), dtype=a.dtype, device=a.device) if top else None
        recv_down = torch.empty((N, h, W, C), dtype=a.dtype, device=a.device) if bottom else None
        if overlap:
            # 1. boundary bands first: their backprops_b reach into the neighbours' halo rows
            part_t = part_b = None
            if top:      # a rows [0, h) meet b_ext rows [0, top + 2h)
                part_t = torch.empty((N, top + 2 * h, W, C), dtype=a.dtype, device=a.device)
                bwd(_rows(gradients, 0, h), _rows(a, 0, h), _rows(b_ext, 0, top + 2 * h), k, md, s1, s2, pad, top, h,
                    precision, ga=_rows(ga, 0, h), gb=part_t)
            if bottom:   # a rows [H - h, H) meet b_ext rows [top + H - 2h, end)
                part_b = torch.empty((N, 2 * h + bottom, W, C), dtype=a.dtype, device=a.device)
                bwd(_rows(gradients, H - h, H), _rows(a, H - h, H), _rows(b_ext, top + H - 2 * h, top + H + bottom), k,
                    md, s1, s2, pad, h, bottom, precision, ga=_rows(ga, H - h, H), gb=part_b)
            # 2. the halo rows of those partial sums go back to their owners ...
            send_up = _rows(part_t, 0, top) if top else None
            send_down = _rows(part_b, 2 * h, 2 * h + bottom) if bottom else None
            reqs = _start_exchange(send_up, recv_up, send_down, recv_down, up, down, group)
            # 3. ... while the interior rows are computed: they touch exactly this rank's own rows of b
            i0, i1 = (h if top else 0), (H - h if bottom else H)
            bwd(_rows(gradients, i0, i1), _rows(a, i0, i1), _rows(b_ext, top, top + H), k, md, s1, s2, pad,
                h if top else 0, h if bottom else 0, precision, ga=_rows(ga, i0, i1), gb=gb)
            # 4. the bands' sums for rows this rank owns
            if top:
                gb[:, :2 * h] += part_t[:, top:]
            if bottom:
                gb[:, H - 2 * h:] += part_b[:, :2 * h]
            _finish(reqs)
        else:
            _, gb_ext = bwd(gradients, a, b_ext, k, md, s1, s2, pad, top, bottom, precision, ga=ga)
            gb.copy_(gb_ext[:, top:top + H])
            send_up = _rows(gb_ext, 0, top) if top else None
            send_down = _rows(gb_ext, top + H, top + H + bottom) if bottom else None
            _finish(_start_exchange(send_up, recv_up, send_down, recv_down, up, down, group))
        # 5. the neighbours' partial sums for this rank's boundary rows
        if top:
            gb[:, :h] += recv_up
        if bottom:
            gb[:, H - h:] += recv_down
        return ga, gb, None, None, None, None


def correlation_strips(a_strip, b_strip, kernel_size, max_displacement, stride_1, stride_2, padding, *, group=None,
                       precision=None, local_ops=None, b_ext=None):
    """Correlation of a row-partitioned pair: every rank passes ITS rows of input_a / input_b and gets its rows of
    the cost volume.  Same positional arguments as ``correlation``.  ``b_ext`` (optional): the tall buffer of
    ``alloc_b_ext`` whose middle rows ARE ``b_strip`` (saves one copy of the strip).  ``local_ops`` (forward, backward
    callables with the signature of ``correlation_halo_forward/backward``) exists for the CPU ``gloo`` tests of the
    exchange logic."""
    args = tuple(int(v) for v in (kernel_size, max_displacement, stride_1, stride_2, padding))
    k, md, s1, s2, pad = args
    if k != 1 or s1 != 1 or pad != md:
        raise ValueError("strip partitioning needs kernel_size 1, stride_1 1 and padding == max_displacement")
    if a_strip.shape != b_strip.shape:
        raise ValueError("input_a and input_b strips must have the same shape")
    if local_ops is None:
        _check_pair(a_strip, b_strip)
        ops = _CUDA_OPS
        precision = _prec(precision, _geom(a_strip, *args))      # resolved once for every band of the strip
    else:
        ops = _cpu_adapter(local_ops)
    if b_ext is not None:
        top, bottom = halo_sizes(group, md, s2)
        N, H, W, C = b_strip.shape
        if tuple(b_ext.shape) != (N, top + H + bottom, W, C) or b_strip.data_ptr() != b_ext[:, top:].data_ptr() or \
                b_strip.stride() != b_ext[:, top:top + H].stride():
            raise ValueError("b_ext must be the tall buffer of alloc_b_ext whose middle rows are b_strip")
    return _CorrelationStrips.apply(a_strip, b_strip, b_ext, args + (precision,), group, ops)
