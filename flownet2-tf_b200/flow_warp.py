"""flow_warp(image, flow) -- drop-in for the reference's ``src/flow_warp.py:7-8`` and its gradient ``:11-15``.

Shapes per ``src/ops/flow_warp/flow_warp_op.cc:6-22``; validation per ``flow_warp.cc:22-31``.
"""
import torch

from . import _lib


def _check(image, flow):
    _lib.require_cuda_f32(image, "image")  # "Input images must have rank 4"
    _lib.require_cuda_f32(flow, "flow")    # "Input flow must have rank 4"
    if image.shape[:3] != flow.shape[:3]:  # flow_warp.cc:26-31
        raise ValueError("Input images and flows must have the same batch size, height and width")
    if flow.shape[3] != 2:
        raise ValueError("flow must have 2 channels")
    if image.device != flow.device:
        raise ValueError("image and flow must be on the same device")


def _forward(image, flow):
    _check(image, flow)
    N, H, W, C = image.shape
    out = torch.empty_like(image)
    with torch.cuda.device(image.device):
        st = _lib.lib().fn2_flow_warp_forward(image.data_ptr(), flow.data_ptr(), out.data_ptr(), N, H, W, C,
                                              _lib.current_stream_ptr(image.device))
    _lib.check(st, "FlowWarp")
    return out


def flow_warp_grad(image, flow, gradients):
    """The ``FlowWarpGrad`` op with both declared outputs (flow_warp_op.cc:13-22): (image_grad, flow_grad)."""
    _check(image, flow)
    _lib.require_cuda_f32(gradients, "gradients")
    if gradients.shape != image.shape:
        raise ValueError("gradients must have the shape of image")
    N, H, W, C = image.shape
    gimage = torch.empty_like(image)  # zero-filled by the library on the same stream
    gflow = torch.empty_like(flow)
    with torch.cuda.device(image.device):
        st = _lib.lib().fn2_flow_warp_backward(image.data_ptr(), flow.data_ptr(), gradients.data_ptr(),
                                               gimage.data_ptr(), gflow.data_ptr(), N, H, W, C,
                                               _lib.current_stream_ptr(image.device))
    _lib.check(st, "FlowWarpGrad")
    return gimage, gflow


class _FlowWarp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, flow):
        ctx.save_for_backward(image, flow)
        return _forward(image, flow)

    @staticmethod
    def backward(ctx, gradients):
        image, flow = ctx.saved_tensors
        return flow_warp_grad(image, flow, gradients.contiguous())


def flow_warp(image, flow):
    return _FlowWarp.apply(image, flow)


def flow_warp_error(image, flow, reference):
    """(warped, brightness_error) = (flow_warp(image, flow), sqrt(sum_c (reference - warped)^2, keepdims)) in one pass
    over the data.  These are the nodes every stacked net runs between two sub-nets
    (``flownet_cs.py:21-27``, ``flownet_css.py:21-27``, ``flownet2.py:33-39``); there they follow frozen nets and data
    inputs, so no gradient exists -- tensors that require grad are refused (use ``flow_warp`` + torch ops instead)."""
    _check(image, flow)
    _lib.require_cuda_f32(reference, "reference")
    if reference.shape != image.shape:
        raise ValueError("reference must have the shape of image")
    if image.requires_grad or flow.requires_grad or reference.requires_grad:
        raise LookupError("flow_warp_error is forward-only (the reference never differentiates these nodes)")
    N, H, W, C = image.shape
    out = torch.empty_like(image)
    err = torch.empty((N, H, W, 1), dtype=torch.float32, device=image.device)
    with torch.cuda.device(image.device):
        st = _lib.lib().fn2_flow_warp_error_forward(image.data_ptr(), flow.data_ptr(), reference.data_ptr(),
                                                    out.data_ptr(), err.data_ptr(), N, H, W, C,
                                                    _lib.current_stream_ptr(image.device))
    _lib.check(st, "FlowWarp")
    return out, err
