"""The caller graphs of the three ops: FlowNetC, S, SD, CS, CSS and FlowNet2.

These are PyTorch restatements of the reference's TF-1/slim graphs -- the harness that runs ``correlation``,
``flow_warp`` and ``downsample`` in the places the reference calls them:

    FlowNetC    src/flownet_c/flownet_c.py:15-125      correlation(conv_a_3, conv_b_3, 1, 20, 1, 2, 20)   (:40)
    FlowNetS    src/flownet_s/flownet_s.py:14-120      6-channel input, or 12 when stacked (:18-26)
    FlowNetSD   src/flownet_sd/flownet_sd.py:14-119    3x3 convs, interconv*, flow = predict_flow2 * 0.05 (:106)
    FlowNetCS   src/flownet_cs/flownet_cs.py:15-38     C (frozen) -> flow_warp -> brightness error -> S
    FlowNetCSS  src/flownet_css/flownet_css.py:15-38   CS (frozen) -> flow_warp -> brightness error -> S
    FlowNet2    src/flownet2/flownet2.py:18-105        CSS || SD -> 2 x flow_warp -> channel norms -> fusion net
    losses      flownet_c.py:127-166, flownet_s.py:122-161, flownet_sd.py:121-160, utils.py:209-224

Tensors at the module boundary are NHWC float32 like the reference's (``inputs['input_a']`` etc.).  Inside, the
convolutions are torch/cuDNN library calls on channels_last tensors, whose memory IS contiguous NHWC: handing a
feature map to one of our NHWC ops is a zero-copy view.  The convolutions are not the hot path of this repo (they are
what TensorFlow's own conv kernels are in the reference); the three custom ops are.

The ops are injected (``ops=``) so that the parity tests can run the same graph, same weights, with a CPU
restatement of the ops standing in for the kernels; the default is this package's CUDA ops, and there is no fallback.
"""
from types import SimpleNamespace

import torch
import torch.nn as nn
import torch.nn.functional as F

LOSS_WEIGHTS = (0.32, 0.08, 0.02, 0.01, 0.005)   # predict_flow6 .. predict_flow2 (flownet_c.py:163)


def default_ops(precision=None):
    """This package's CUDA ops behind the reference's three signatures (+ the fused multi-scale downsample)."""
    from .correlation import correlation as _correlation, correlation_leaky_concat as _fused
    from .flow_warp import flow_warp as _flow_warp, flow_warp_error as _flow_warp_error
    from .downsample import downsample as _downsample, downsample_multi as _downsample_multi, \
        multiscale_epe as _multiscale_epe

    def corr(a, b, k, md, s1, s2, pad):
        return _correlation(a, b, k, md, s1, s2, pad, precision=precision)

    def corr_fused(a, b, k, md, s1, s2, pad, buf, offset, slope):
        return _fused(a, b, k, md, s1, s2, pad, buf, offset, slope, precision=precision)
    return SimpleNamespace(correlation=corr, flow_warp=_flow_warp, downsample=_downsample,
                           downsample_multi=_downsample_multi, multiscale_epe=_multiscale_epe,
                           correlation_leaky_concat=corr_fused,
                           flow_warp_error=_flow_warp_error)


def leaky_relu(x):
    # utils.py:401-405: 0.55 x + 0.45 |x|, i.e. slope 0.1 for x < 0
    return F.leaky_relu(x, 0.1)


def nhwc(x):
    """NCHW tensor (channels_last memory) -> contiguous NHWC view for the ops."""
    return x.permute(0, 2, 3, 1).contiguous()


def nchw(x):
    """NHWC tensor -> NCHW view with channels_last memory (no copy)."""
    return x.permute(0, 3, 1, 2)


def channel_norm(x):
    """sqrt(sum over channels of x^2), NCHW, keepdim (flownet2.py:25-28, flownet_cs.py:24-27)."""
    return torch.sqrt(torch.sum(x * x, dim=1, keepdim=True))


class Conv(nn.Module):
    """pad(x, (k-1)/2) + slim.conv2d(..., padding='VALID') [+ LeakyReLU] (flownet_c.py:18-31, utils.py:408-412)."""

    def __init__(self, cin, cout, k, stride=1, act=True):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, k, stride=stride, padding=(k - 1) // 2, bias=True)
        self.act = act

    def forward(self, x):
        y = self.conv(x)
        return leaky_relu(y) if self.act else y


class Deconv(nn.Module):
    """antipad(slim.conv2d_transpose(x, cout, 4, stride=2, padding='VALID')) [+ LeakyReLU]: the VALID transposed
    convolution yields 2h+2 rows, antipad crops one on each side (utils.py:415-421) = ConvTranspose2d(4, 2, 1)."""

    def __init__(self, cin, cout, act=True, bias=False):
        super().__init__()
        self.deconv = nn.ConvTranspose2d(cin, cout, 4, stride=2, padding=1, bias=bias)
        self.act = act

    def forward(self, x):
        y = self.deconv(x)
        return leaky_relu(y) if self.act else y


def init_like_reference(module):
    """slim.variance_scaling_initializer() (He, fan-in) for weights, zeros for biases (flownet_c.py:21)."""
    for m in module.modules():
        if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
            nn.init.kaiming_normal_(m.weight, a=0.0, mode="fan_in", nonlinearity="relu")
            if m.bias is not None:
                nn.init.zeros_(m.bias)
    return module


class Refinement(nn.Module):
    """The decoder shared by FlowNetC and FlowNetS (flownet_c.py:56-108, flownet_s.py:52-104): predict_flow6..2 (3x3,
    no activation), deconv5..2 (512, 256, 128, 64; no bias), upsample_flow* (2 channels, no bias, no activation),
    skip concatenations [conv_k, deconv_k, upsampled flow].  ``inter`` adds FlowNetSD's interconv layers
    (flownet_sd.py:56-101), which smooth the concatenation before each prediction."""

    def __init__(self, c6=1024, c5=512, c4=512, c3=256, c2=128, inter=False):
        super().__init__()
        self.predict_flow6 = Conv(c6, 2, 3, act=False)
        self.deconv5 = Deconv(c6, 512)
        self.upsample_flow6to5 = Deconv(2, 2, act=False)
        cat5 = c5 + 512 + 2
        self.deconv4 = Deconv(cat5, 256)
        self.upsample_flow5to4 = Deconv(2, 2, act=False)
        cat4 = c4 + 256 + 2
        self.deconv3 = Deconv(cat4, 128)
        self.upsample_flow4to3 = Deconv(2, 2, act=False)
        cat3 = c3 + 128 + 2
        self.deconv2 = Deconv(cat3, 64)
        self.upsample_flow3to2 = Deconv(2, 2, act=False)
        cat2 = c2 + 64 + 2
        self.inter = inter
        if inter:
            self.interconv5 = Conv(cat5, 512, 3, act=False)
            self.interconv4 = Conv(cat4, 256, 3, act=False)
            self.interconv3 = Conv(cat3, 128, 3, act=False)
            self.interconv2 = Conv(cat2, 64, 3, act=False)
            cat5, cat4, cat3, cat2 = 512, 256, 128, 64
        self.predict_flow5 = Conv(cat5, 2, 3, act=False)
        self.predict_flow4 = Conv(cat4, 2, 3, act=False)
        self.predict_flow3 = Conv(cat3, 2, 3, act=False)
        self.predict_flow2 = Conv(cat2, 2, 3, act=False)

    def forward(self, conv6_1, conv5_1, conv4_1, conv3_1, conv2):
        pf6 = self.predict_flow6(conv6_1)
        concat5 = torch.cat([conv5_1, self.deconv5(conv6_1), self.upsample_flow6to5(pf6)], dim=1)
        pf5 = self.predict_flow5(self.interconv5(concat5) if self.inter else concat5)
        concat4 = torch.cat([conv4_1, self.deconv4(concat5), self.upsample_flow5to4(pf5)], dim=1)
        pf4 = self.predict_flow4(self.interconv4(concat4) if self.inter else concat4)
        concat3 = torch.cat([conv3_1, self.deconv3(concat4), self.upsample_flow4to3(pf4)], dim=1)
        pf3 = self.predict_flow3(self.interconv3(concat3) if self.inter else concat3)
        concat2 = torch.cat([conv2, self.deconv2(concat3), self.upsample_flow3to2(pf3)], dim=1)
        pf2 = self.predict_flow2(self.interconv2(concat2) if self.inter else concat2)
        return pf6, pf5, pf4, pf3, pf2


def _predictions(pfs, flow_scale, height, width):
    """The reference's output dict (flownet_c.py:110-125), NHWC: predict_flow6..2 and the full-resolution flow =
    predict_flow2 * flow_scale resized bilinearly with align_corners=True."""
    pf6, pf5, pf4, pf3, pf2 = pfs
    flow = F.interpolate(pf2 * flow_scale, size=(height, width), mode="bilinear", align_corners=True)
    out = {"predict_flow%d" % k: nhwc(p) for k, p in zip((6, 5, 4, 3, 2), pfs)}
    out["flow"] = nhwc(flow)
    return out


def _to_cl(x_nhwc):
    return nchw(x_nhwc).contiguous(memory_format=torch.channels_last)


class FlowNetC(nn.Module):
    """flownet_c.py:15-125."""

    def __init__(self, ops=None, precision=None):
        super().__init__()
        self.ops = ops if ops is not None else default_ops(precision)
        self.conv1 = Conv(3, 64, 7, 2)
        self.conv2 = Conv(64, 128, 5, 2)
        self.conv3 = Conv(128, 256, 5, 2)
        self.conv_redir = Conv(256, 32, 1, 1)
        self.conv3_1 = Conv(473, 256, 3)
        self.conv4 = Conv(256, 512, 3, 2)
        self.conv4_1 = Conv(512, 512, 3)
        self.conv5 = Conv(512, 512, 3, 2)
        self.conv5_1 = Conv(512, 512, 3)
        self.conv6 = Conv(512, 1024, 3, 2)
        self.conv6_1 = Conv(1024, 1024, 3)
        self.refine = Refinement()
        init_like_reference(self)

    def forward(self, inputs):
        a, b = inputs["input_a"], inputs["input_b"]
        height, width = a.shape[1], a.shape[2]
        n = a.shape[0]
        # the two towers share their weights (reuse=True, :34-37): run them as one batch of 2N
        x = _to_cl(torch.cat([a, b], dim=0))
        c1 = self.conv1(x)
        c2 = self.conv2(c1)
        c3 = self.conv3(c2)
        conv_a_2, conv_a_3, conv_b_3 = c2[:n], c3[:n], c3[n:]
        redir = self.conv_redir(conv_a_3)                                            # :44
        if hasattr(self.ops, "correlation_leaky_concat"):
            # correlation + LeakyReLU + concat (:40-46) in the kernel's epilogue: the 441 channels land directly in
            # the 473-channel buffer behind conv_redir's 32
            buf = torch.empty((n, c3.shape[2], c3.shape[3], 32 + 441), dtype=c3.dtype, device=c3.device)
            buf[..., :32] = redir.permute(0, 2, 3, 1)
            buf = self.ops.correlation_leaky_concat(nhwc(conv_a_3), nhwc(conv_b_3), 1, 20, 1, 2, 20, buf, 32, 0.1)
            net = nchw(buf)
        else:
            cc = self.ops.correlation(nhwc(conv_a_3), nhwc(conv_b_3), 1, 20, 1, 2, 20)   # :40
            cc_relu = leaky_relu(nchw(cc))                                               # :41
            net = torch.cat([redir, cc_relu], dim=1)                                     # :44-46, 32 + 441 channels
        conv3_1 = self.conv3_1(net)
        conv4_1 = self.conv4_1(self.conv4(conv3_1))
        conv5_1 = self.conv5_1(self.conv5(conv4_1))
        conv6_1 = self.conv6_1(self.conv6(conv5_1))
        pfs = self.refine(conv6_1, conv5_1, conv4_1, conv3_1, conv_a_2)              # skip from conv_a_2 (:105)
        return _predictions(pfs, 20.0, height, width)


class FlowNetS(nn.Module):
    """flownet_s.py:14-120.  ``stacked`` selects the 12-channel input [a, b, warped, flow, brightness_error]."""

    def __init__(self, stacked=False):
        super().__init__()
        self.stacked = stacked
        self.conv1 = Conv(12 if stacked else 6, 64, 7, 2)
        self.conv2 = Conv(64, 128, 5, 2)
        self.conv3 = Conv(128, 256, 5, 2)
        self.conv3_1 = Conv(256, 256, 3)
        self.conv4 = Conv(256, 512, 3, 2)
        self.conv4_1 = Conv(512, 512, 3)
        self.conv5 = Conv(512, 512, 3, 2)
        self.conv5_1 = Conv(512, 512, 3)
        self.conv6 = Conv(512, 1024, 3, 2)
        self.conv6_1 = Conv(1024, 1024, 3)
        self.refine = Refinement()
        init_like_reference(self)

    def forward(self, inputs):
        a = inputs["input_a"]
        height, width = a.shape[1], a.shape[2]
        if "warped" in inputs and "flow" in inputs and "brightness_error" in inputs:   # :18-24
            parts = [a, inputs["input_b"], inputs["warped"], inputs["flow"], inputs["brightness_error"]]
        else:
            parts = [a, inputs["input_b"]]
        x = _to_cl(torch.cat(parts, dim=3))
        conv2 = self.conv2(self.conv1(x))
        conv3_1 = self.conv3_1(self.conv3(conv2))
        conv4_1 = self.conv4_1(self.conv4(conv3_1))
        conv5_1 = self.conv5_1(self.conv5(conv4_1))
        conv6_1 = self.conv6_1(self.conv6(conv5_1))
        pfs = self.refine(conv6_1, conv5_1, conv4_1, conv3_1, conv2)
        return _predictions(pfs, 20.0, height, width)


class FlowNetSD(nn.Module):
    """flownet_sd.py:14-119: small displacements; 3x3 convolutions, interconv layers, flow = predict_flow2 * 0.05."""

    def __init__(self):
        super().__init__()
        self.conv0 = Conv(6, 64, 3)
        self.conv1 = Conv(64, 64, 3, 2)
        self.conv1_1 = Conv(64, 128, 3)
        self.conv2 = Conv(128, 128, 3, 2)
        self.conv2_1 = Conv(128, 128, 3)
        self.conv3 = Conv(128, 256, 3, 2)
        self.conv3_1 = Conv(256, 256, 3)
        self.conv4 = Conv(256, 512, 3, 2)
        self.conv4_1 = Conv(512, 512, 3)
        self.conv5 = Conv(512, 512, 3, 2)
        self.conv5_1 = Conv(512, 512, 3)
        self.conv6 = Conv(512, 1024, 3, 2)
        self.conv6_1 = Conv(1024, 1024, 3)
        self.refine = Refinement(inter=True)
        init_like_reference(self)

    def forward(self, inputs):
        a = inputs["input_a"]
        height, width = a.shape[1], a.shape[2]
        x = _to_cl(torch.cat([a, inputs["input_b"]], dim=3))
        conv2 = self.conv2(self.conv1_1(self.conv1(self.conv0(x))))
        conv3_1 = self.conv3_1(self.conv3(self.conv2_1(conv2)))
        conv4_1 = self.conv4_1(self.conv4(conv3_1))
        conv5_1 = self.conv5_1(self.conv5(conv4_1))
        conv6_1 = self.conv6_1(self.conv6(conv5_1))
        pfs = self.refine(conv6_1, conv5_1, conv4_1, conv3_1, conv2)     # skip from conv2, not conv2_1 (:97)
        return _predictions(pfs, 0.05, height, width)


def _warp_and_error(ops, inputs, flow_nhwc):
    """flow_warp(input_b, flow) and the brightness error ||input_a - warped||_2 over channels
    (flownet_cs.py:21-27, flownet_css.py:21-27), both NHWC."""
    a, b = inputs["input_a"], inputs["input_b"]
    if hasattr(ops, "flow_warp_error") and not (a.requires_grad or b.requires_grad or flow_nhwc.requires_grad):
        return ops.flow_warp_error(b.contiguous(), flow_nhwc.contiguous(), a.contiguous())   # one pass over the data
    warped = ops.flow_warp(inputs["input_b"].contiguous(), flow_nhwc.contiguous())
    diff = inputs["input_a"] - warped
    return warped, torch.sqrt(torch.sum(diff * diff, dim=3, keepdim=True))


class FlowNetCS(nn.Module):
    """flownet_cs.py:15-38: FlowNetC with frozen weights (trainable=False, :18), then FlowNetS on
    [a, b, warp(b, flow_C), flow_C * 0.05, brightness error]."""

    def __init__(self, ops=None, precision=None):
        super().__init__()
        self.ops = ops if ops is not None else default_ops(precision)
        self.net_c = FlowNetC(self.ops)
        self.net_s = FlowNetS(stacked=True)

    def forward(self, inputs):
        with torch.no_grad():
            flow_c = self.net_c(inputs)["flow"]
        warped, err = _warp_and_error(self.ops, inputs, flow_c)
        return self.net_s({"input_a": inputs["input_a"], "input_b": inputs["input_b"], "warped": warped,
                           "flow": flow_c * 0.05, "brightness_error": err})


class FlowNetCSS(nn.Module):
    """flownet_css.py:15-38: FlowNetCS frozen, then a second FlowNetS."""

    def __init__(self, ops=None, precision=None):
        super().__init__()
        self.ops = ops if ops is not None else default_ops(precision)
        self.net_cs = FlowNetCS(self.ops)
        self.net_s = FlowNetS(stacked=True)

    def forward(self, inputs):
        with torch.no_grad():
            flow_cs = self.net_cs(inputs)["flow"]
        warped, err = _warp_and_error(self.ops, inputs, flow_cs)
        return self.net_s({"input_a": inputs["input_a"], "input_b": inputs["input_b"], "warped": warped,
                           "flow": flow_cs * 0.05, "brightness_error": err})


class FlowNet2(nn.Module):
    """flownet2.py:18-105: FlowNetCSS and FlowNetSD (both frozen, :22-23), two flow_warps, channel norms of both flows
    and both image differences, 11-channel input to the fusion net, whose transposed convolutions DO have biases
    (no biases_initializer=None scope there, :50-98); predict_flow0 at full resolution."""

    def __init__(self, ops=None, precision=None):
        super().__init__()
        self.ops = ops if ops is not None else default_ops(precision)
        self.net_css = FlowNetCSS(self.ops)
        self.net_sd = FlowNetSD()
        self.fuse_conv0 = Conv(11, 64, 3)
        self.fuse_conv1 = Conv(64, 64, 3, 2)
        self.fuse_conv1_1 = Conv(64, 128, 3)
        self.fuse_conv2 = Conv(128, 128, 3, 2)
        self.fuse_conv2_1 = Conv(128, 128, 3)
        self.predict_flow2 = Conv(128, 2, 3, act=False)
        self.fuse_deconv1 = Deconv(128, 32, bias=True)
        self.fuse_upsample_flow2to1 = Deconv(2, 2, act=False, bias=True)
        self.fuse_interconv1 = Conv(128 + 32 + 2, 32, 3, act=False)
        self.predict_flow1 = Conv(32, 2, 3, act=False)
        self.fuse_deconv0 = Deconv(128 + 32 + 2, 16, bias=True)
        self.fuse_upsample_flow1to0 = Deconv(2, 2, act=False, bias=True)
        self.fuse_interconv0 = Conv(64 + 16 + 2, 16, 3, act=False)
        self.predict_flow0 = Conv(16, 2, 3, act=False)
        for m in (self.fuse_conv0, self.fuse_conv1, self.fuse_conv1_1, self.fuse_conv2, self.fuse_conv2_1,
                  self.predict_flow2, self.fuse_deconv1, self.fuse_upsample_flow2to1, self.fuse_interconv1,
                  self.predict_flow1, self.fuse_deconv0, self.fuse_upsample_flow1to0, self.fuse_interconv0,
                  self.predict_flow0):
            init_like_reference(m)

    def forward(self, inputs):
        a = inputs["input_a"]
        height, width = a.shape[1], a.shape[2]
        with torch.no_grad():
            flow_css = self.net_css(inputs)["flow"]
            flow_sd = self.net_sd(inputs)["flow"]
        norm = lambda t: torch.sqrt(torch.sum(t * t, dim=3, keepdim=True))   # ChannelNorm, NHWC (:25-28)
        warp_sd, diff_sd_norm = _warp_and_error(self.ops, inputs, flow_sd)    # :33-35
        warp_css, diff_css_norm = _warp_and_error(self.ops, inputs, flow_css)  # :37-39
        x = _to_cl(torch.cat([a, flow_sd, flow_css, norm(flow_sd), norm(flow_css), diff_sd_norm, diff_css_norm],
                             dim=3))                                          # :41-47, 3+2+2+1+1+1+1 = 11 channels
        conv0 = self.fuse_conv0(x)
        conv1_1 = self.fuse_conv1_1(self.fuse_conv1(conv0))
        conv2_1 = self.fuse_conv2_1(self.fuse_conv2(conv1_1))
        pf2 = self.predict_flow2(conv2_1)
        concat1 = torch.cat([conv1_1, self.fuse_deconv1(conv2_1), self.fuse_upsample_flow2to1(pf2)], dim=1)
        pf1 = self.predict_flow1(self.fuse_interconv1(concat1))
        concat0 = torch.cat([conv0, self.fuse_deconv0(concat1), self.fuse_upsample_flow1to0(pf1)], dim=1)
        pf0 = self.predict_flow0(self.fuse_interconv0(concat0))
        flow = F.interpolate(pf0, size=(height, width), mode="bilinear", align_corners=True)
        return {"predict_flow0": nhwc(pf0), "flow": nhwc(flow)}


# ---------------------------------------------------------------------------------------------------- losses
def average_endpoint_error(labels, predictions):
    """utils.py:209-224: sum over pixels of ||labels - predictions||_2 (over the 2 flow channels), divided by N."""
    d = predictions - labels
    return torch.sqrt(torch.sum(d * d, dim=3)).sum() / predictions.shape[0]


def weighted_loss(losses, weights):
    """tf.losses.compute_weighted_loss(losses, weights) with its default reduction SUM_BY_NONZERO_WEIGHTS
    (flownet_c.py:163): sum_i w_i * l_i divided by the number of non-zero weights -- 5 for the FlowNet losses, so
    the data term is a fifth of the plain weighted sum relative to the L2 regulariser."""
    nonzero = sum(1 for w in weights if w != 0)
    total = sum(w * l for w, l in zip(weights, losses))
    return total / nonzero if nonzero else total * 0.0


def multiscale_loss(flow, predictions, ops, gt_scale=0.05, weights=LOSS_WEIGHTS):
    """The training loss of FlowNetC / S (gt_scale 0.05: flownet_c.py:127-166, flownet_s.py:122-161) and SD
    (gt_scale 20.0: flownet_sd.py:121-160): the ground-truth flow, scaled, is downsampled to the size of each of
    predict_flow6..2 and compared by average end-point error; weights 0.32, 0.08, 0.02, 0.01, 0.005.  The five
    downsample() calls of the reference are one fused pass here (``ops.downsample_multi``) when the ops offer it."""
    preds = [predictions["predict_flow%d" % k] for k in (6, 5, 4, 3, 2)]
    if hasattr(ops, "multiscale_epe"):
        # one fused pass: flow * gt_scale on load, the five downsamples, the five end-point-error sums (f-3)
        epe = ops.multiscale_epe(flow.contiguous(), [p.contiguous() for p in preds], gt_scale)
        losses = [epe[i] for i in range(len(preds))]
        return weighted_loss(losses, weights), losses
    flow = (flow * gt_scale).contiguous()
    sizes = [(int(p.shape[1]), int(p.shape[2])) for p in preds]
    if hasattr(ops, "downsample_multi"):
        downs = ops.downsample_multi(flow, sizes)
    else:
        downs = [ops.downsample(flow, list(s)) for s in sizes]
    losses = [average_endpoint_error(d, p) for d, p in zip(downs, preds)]
    return weighted_loss(losses, weights), losses


def average_endpoint_error_hfem(labels, predictions, add_hfem="", lambda_w=2.0, perc_hfem=50, edges=None):
    """utils.py:227-339.  '' (or 'edges' without an edge map): the plain average end-point error.  'edges': every
    pixel's error counts 1 + lambda_w * edges times (:326-334).  'hard': only the top perc_hfem per cent of the errors
    count, (1 + lambda_w) times, rescaled by total / hard pixels (:268-324; the selection itself carries no
    gradient, as tf.nn.top_k's indices)."""
    n = predictions.shape[0]
    d = predictions - labels
    epe = torch.sqrt(torch.sum(d * d, dim=3, keepdim=True))
    mode = (add_hfem or "").lower()
    if mode == "hard":
        flat = epe.reshape(-1)
        k = int(round(float(perc_hfem) / 100.0 * flat.numel()))     # tf.round: half to even, as Python's round
        top = torch.topk(flat, k).values
        return (1.0 + lambda_w) * top.sum() / n * (flat.numel() / float(k))
    if mode == "edges" and edges is not None:
        return (epe + lambda_w * (epe * edges)).sum() / n
    return epe.sum() / n


def multiscale_loss_hfem(flow, predictions, ops, gt_scale=0.05, weights=LOSS_WEIGHTS, add_hfem="", lambda_w=2.0,
                         perc_hfem=50, edges=None):
    """The loss of the reference's HFEM variants (flownet_s_interp.py:160-245 and the other *_interp nets): as
    ``multiscale_loss`` with ``average_endpoint_error_hfem`` per scale; in mode 'edges' the edge map [N,H,W,1] is
    downsampled to every scale as well (:177,191,205,219,233).  With the library's ops the 'edges' mode is two fused
    passes: all scales of the edge map, then flow scaling + downsampling + weighted error sums
    (``ops.multiscale_epe(..., edges=)``, fn2_multiscale_epe_edges_forward)."""
    preds = [predictions["predict_flow%d" % k] for k in (6, 5, 4, 3, 2)]
    mode = (add_hfem or "").lower()
    if mode not in ("hard", "edges") or (mode == "edges" and edges is None):
        return multiscale_loss(flow, predictions, ops, gt_scale, weights)
    if mode == "edges" and hasattr(ops, "multiscale_epe"):
        epe = ops.multiscale_epe(flow.contiguous(), [p.contiguous() for p in preds], gt_scale,
                                 edges=edges.contiguous(), lambda_w=lambda_w)
        losses = [epe[i] for i in range(len(preds))]
        return weighted_loss(losses, weights), losses
    sizes = [(int(p.shape[1]), int(p.shape[2])) for p in preds]
    many = getattr(ops, "downsample_multi", lambda x, ss: [ops.downsample(x, list(s)) for s in ss])
    downs = many((flow * gt_scale).contiguous(), sizes)
    emaps = many(edges.contiguous(), sizes) if mode == "edges" else [None] * len(preds)
    losses = [average_endpoint_error_hfem(d, p, mode, lambda_w, perc_hfem, e) for d, p, e in zip(downs, preds, emaps)]
    return weighted_loss(losses, weights), losses


def l2_regularization(module, scale):
    """slim.l2_regularizer(scale) on the conv2d weights only (flownet_c.py:27-28; transposed convolutions and biases
    carry none): scale * sum(w^2) / 2, added by tf.losses.get_total_loss() (:166)."""
    total = 0.0
    for m in module.modules():
        if isinstance(m, nn.Conv2d) and m.weight.requires_grad:
            total = total + 0.5 * scale * torch.sum(m.weight * m.weight)
    return total
