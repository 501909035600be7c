"""downsample(tensor, size) -- drop-in for the reference's ``src/downsample.py:7-8``.

``size`` = [out_height, out_width] (the op attr, ``src/ops/downsample/downsample_op.cc:24-28``).  The reference
registers no gradient for this op; differentiating through it raises LookupError here as tf.gradients does there.
"""
import torch

from . import _lib


def _forward(tensor, size):
    _lib.require_cuda_f32(tensor, "input")  # "Input images must have rank 4" (downsample_kernel.cc:25)
    size = list(size)
    if len(size) != 2:  # downsample_kernel.cc:19
        raise ValueError("size must be 2 dimensions")
    oh, ow = int(size[0]), int(size[1])
    N, H, W, C = tensor.shape
    if oh < 1 or ow < 1:
        raise ValueError("Downsample: output dimensions must be positive")
    out = torch.empty((N, oh, ow, C), dtype=torch.float32, device=tensor.device)
    with torch.cuda.device(tensor.device):
        st = _lib.lib().fn2_downsample_forward(tensor.data_ptr(), out.data_ptr(), N, H, W, C, oh, ow,
                                               _lib.current_stream_ptr(tensor.device))
    _lib.check(st, "Downsample")
    return out


class _Downsample(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tensor, oh, ow):
        return _forward(tensor, (oh, ow))

    @staticmethod
    def backward(ctx, gradients):
        raise LookupError("No gradient defined for operation 'Downsample' (the reference registers none)")


def downsample(tensor, size):
    size = list(size)
    if len(size) != 2:
        raise ValueError("size must be 2 dimensions")
    if isinstance(tensor, torch.Tensor) and tensor.requires_grad:
        return _Downsample.apply(tensor, int(size[0]), int(size[1]))
    return _forward(tensor, size)


def _check_sizes(sizes):
    sizes = [list(s) for s in sizes]
    for s in sizes:
        if len(s) != 2:  # downsample_kernel.cc:19
            raise ValueError("size must be 2 dimensions")
        if int(s[0]) < 1 or int(s[1]) < 1:
            raise ValueError("Downsample: output dimensions must be positive")
    return [(int(s[0]), int(s[1])) for s in sizes]


def _multi_workspace(tensor, hw, nscales):
    """Scratch for the fused multi-scale kernels (weight tables + per-tile partial sums); the library allocates nothing."""
    import ctypes
    N, H, W, C = tensor.shape
    nbytes = _lib.lib().fn2_downsample_multi_workspace_bytes(hw, nscales, N, H, W, C)
    if nbytes == 0:
        return None, 0, ctypes.c_void_p(0)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=tensor.device)
    return ws, nbytes, ctypes.c_void_p(ws.data_ptr())


def downsample_multi(tensor, sizes):
    """[downsample(tensor, size) for size in sizes] in one pass over ``tensor``.

    The reference's losses downsample the same ground-truth flow to five sizes with five op calls
    (``flownet_c.py:135-160``); each output here equals the single-size op's.  No gradient, as for ``downsample``.
    """
    import ctypes
    _lib.require_cuda_f32(tensor, "input")
    if isinstance(tensor, torch.Tensor) and tensor.requires_grad:
        raise LookupError("No gradient defined for operation 'Downsample' (the reference registers none)")
    sizes = _check_sizes(sizes)
    if not sizes:
        return []
    N, H, W, C = tensor.shape
    outs = [torch.empty((N, h, w, C), dtype=torch.float32, device=tensor.device) for h, w in sizes]
    ptrs = (ctypes.c_void_p * len(outs))(*[o.data_ptr() for o in outs])
    hw = (ctypes.c_int * (2 * len(outs)))(*[v for s in sizes for v in s])
    with torch.cuda.device(tensor.device):
        ws, nbytes, wsp = _multi_workspace(tensor, hw, len(outs))
        st = _lib.lib().fn2_downsample_multi_forward(tensor.data_ptr(), ptrs, hw, len(outs), N, H, W, C, wsp, nbytes,
                                                     _lib.current_stream_ptr(tensor.device))
    _lib.check(st, "Downsample")
    return outs


class _MultiscaleEpe(torch.autograd.Function):
    """epe[s] = average_endpoint_error(downsample(flow * scale, size_s), predictions[s]) for every s, one pass.

    Differentiable with respect to the predictions (the kernel also writes (pred - down) / epe, which is all the
    gradient needs); the ground-truth flow gets none, as in the reference (downsample has no registered gradient)."""

    @staticmethod
    def forward(ctx, flow, scale, want_down, edge_maps, lambda_w, *preds):
        import ctypes
        N, H, W, C = flow.shape
        sizes = [(int(p.shape[1]), int(p.shape[2])) for p in preds]
        k = len(preds)
        need_grad = any(ctx.needs_input_grad[5:])
        dirs = [torch.empty_like(p) for p in preds] if need_grad else []
        downs = [torch.empty_like(p) for p in preds] if want_down else []
        epe = torch.empty(k, dtype=torch.float32, device=flow.device)
        arr = ctypes.c_void_p * k
        pp = arr(*[p.data_ptr() for p in preds])
        dp = arr(*[d.data_ptr() for d in dirs]) if dirs else None
        op = arr(*[d.data_ptr() for d in downs]) if downs else None
        hw = (ctypes.c_int * (2 * k))(*[v for s in sizes for v in s])
        with torch.cuda.device(flow.device):
            ws, nbytes, wsp = _multi_workspace(flow, hw, k)
            if edge_maps is None:
                st = _lib.lib().fn2_multiscale_epe_forward(flow.data_ptr(), float(scale), pp, hw, k, N, H, W, C, op, dp,
                                                           epe.data_ptr(), wsp, nbytes,
                                                           _lib.current_stream_ptr(flow.device))
            else:
                ep = arr(*[e.data_ptr() for e in edge_maps])
                st = _lib.lib().fn2_multiscale_epe_edges_forward(flow.data_ptr(), float(scale), pp, ep, float(lambda_w),
                                                                 hw, k, N, H, W, C, op, dp, epe.data_ptr(), wsp, nbytes,
                                                                 _lib.current_stream_ptr(flow.device))
        _lib.check(st, "MultiscaleEpe")
        ctx.save_for_backward(*dirs)
        ctx.n, ctx.k = N, k
        ctx.mark_non_differentiable(*downs)
        return (epe,) + tuple(downs)

    @staticmethod
    def backward(ctx, g_epe, *unused):
        dirs = ctx.saved_tensors
        if not dirs:
            return (None,) * (5 + ctx.k)
        grads = tuple((g_epe[s] / ctx.n) * d if ctx.needs_input_grad[5 + s] else None for s, d in enumerate(dirs))
        return (None,) * 5 + grads


def multiscale_epe(flow, predictions, flow_scale=1.0, return_downsampled=False, edges=None, lambda_w=2.0):
    """Per-scale data terms of the FlowNet training loss in ONE pass over ``flow``.

    For predictions p_s of shape [N,h_s,w_s,C] returns the tensor
        epe[s] = sum_pixels sqrt(sum_c (p_s - downsample(flow * flow_scale, [h_s, w_s]))^2) / N
    i.e. ``average_endpoint_error(downsample(flow * 0.05, size), predict_flow)`` of ``flownet_c.py:127-160`` /
    ``utils.py:209-224`` for all scales, without materialising the scaled flow or re-reading it per scale, and with the
    reductions done by the kernel in a fixed order.  Differentiable w.r.t. the predictions.  With
    ``return_downsampled=True`` also returns the list of downsampled tensors.

    ``edges`` ([N,H,W,1], values in [0,1]) switches on the edge weighting of the reference's HFEM loss in mode 'edges'
    (``average_endpoint_error_hfem``, ``utils.py:326-334``; ``flownet_s_interp.py:170-239``): the map is downsampled
    to every scale in one fused pass of its own and every pixel's error then counts ``1 + lambda_w * edges_s`` times,
        epe[s] = sum_pixels (epe_s + lambda_w * epe_s * downsample(edges, size_s)) / N.

    Shapes outside the fused kernels' limits (``fn2_multiscale_epe_supported``: C <= 2, W * C a multiple of 4 and
    <= 1580, reductions by >= 1.5) get the same values from ``downsample_multi`` on the scaled flow plus the
    end-point-error nodes in torch: slower (the flow is scaled in a pass of its own and the sums are torch's), same
    result within the op's tolerance."""
    _lib.require_cuda_f32(flow, "flow")
    if flow.requires_grad:
        raise LookupError("No gradient defined for operation 'Downsample' (the reference registers none)")
    predictions = list(predictions)
    if not predictions:
        raise ValueError("at least one prediction is needed")
    N, H, W, C = flow.shape
    for p in predictions:
        _lib.require_cuda_f32(p, "prediction")
        if p.shape[0] != N or p.shape[3] != C:
            raise ValueError("prediction %s does not match flow %s" % (tuple(p.shape), tuple(flow.shape)))
        if p.device != flow.device:
            raise ValueError("predictions and flow must be on the same device")
    import ctypes
    sizes = [(int(p.shape[1]), int(p.shape[2])) for p in predictions]
    edge_maps = None
    if edges is not None:
        _lib.require_cuda_f32(edges, "edges")
        if tuple(edges.shape) != (N, H, W, 1) or edges.device != flow.device:
            raise ValueError("edges must be [N,H,W,1] on the flow's device, got %s" % (tuple(edges.shape),))
        if edges.requires_grad:
            raise LookupError("No gradient defined for operation 'Downsample' (the reference registers none)")
        edge_maps = downsample_multi(edges, sizes)          # flownet_s_interp.py:177,191,205,219,233 in one pass
    hw = (ctypes.c_int * (2 * len(sizes)))(*[v for s in sizes for v in s])
    fused = _lib.lib().fn2_multiscale_epe_supported(hw, len(sizes), N, H, W, C)
    if fused < 0:
        _lib.check(fused, "MultiscaleEpe")
    if fused == 0:
        # Shapes the fused kernels do not take (flows wider than 790 px at C = 2 -- Sintel, KITTI, 4K --, widths with
        # W * C % 4 != 0, C > 2, reductions below 1.5): the reference's own node sequence on this library's
        # downsample kernels -- flow * scale materialised, one downsample_multi call (which itself goes scale by scale
        # where it must), average_endpoint_error by torch autograd.
        downs = downsample_multi((flow * float(flow_scale)).contiguous(), sizes)
        maps = [torch.sqrt(torch.sum((p - d) * (p - d), dim=3, keepdim=True)) for p, d in zip(predictions, downs)]
        if edge_maps is not None:
            maps = [m + float(lambda_w) * (m * e) for m, e in zip(maps, edge_maps)]
        epe = torch.stack([m.sum() / N for m in maps])
        return (epe, downs) if return_downsampled else epe
    res = _MultiscaleEpe.apply(flow, float(flow_scale), bool(return_downsampled), edge_maps, float(lambda_w),
                               *predictions)
    return (res[0], list(res[1:])) if return_downsampled else res[0]
