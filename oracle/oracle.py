"""Python face of the oracle.  TEST INFRASTRUCTURE ONLY (never imported by the product path).

Two independent CPU restatements of the reference op semantics:
  * ``c_*``  -- ctypes calls into oracle/libfn2oracle.so (fn2_oracle.c: plain C loops, order-faithful fp32 and
                double-accumulation variants, OpenMP over rows; also the timed CPU baseline);
  * ``np_*`` -- vectorised numpy in float64, written from the formulas rather than from the kernels' loop
                structure (padded arrays + shifts for correlation, explicit adjoint for its gradient).
tests/test_oracle.py requires the two to agree; tests/test_reference_pin.py pins them to the reference's own CUDA
kernels (oracle/_ref).  Reference citations are in fn2_oracle.c.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libfn2oracle.so")
_lib = None

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_i = ctypes.c_int


def build():
    res = subprocess.run(["make", "-C", _HERE, "libfn2oracle.so"], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + res.stdout + res.stderr)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = ctypes.CDLL(_SO)
        L.fn2o_correlation_output_shape.argtypes = [_i] * 7 + [ctypes.POINTER(_i)] * 3
        L.fn2o_correlation_forward.argtypes = [_f32p, _f32p, _f32p] + [_i] * 9
        L.fn2o_correlation_forward_f64.argtypes = [_f32p, _f32p, _f64p] + [_i] * 9
        L.fn2o_correlation_backward.argtypes = [_f32p] * 5 + [_i] * 9
        L.fn2o_correlation_backward_f64.argtypes = [_f32p] * 3 + [_f64p] * 2 + [_i] * 9
        L.fn2o_flow_warp_forward.argtypes = [_f32p] * 3 + [_i] * 4
        L.fn2o_flow_warp_backward.argtypes = [_f32p] * 5 + [_i] * 4
        L.fn2o_downsample_forward.argtypes = [_f32p] * 2 + [_i] * 6
        _lib = L
    return _lib


def _c(x, dt=np.float32):
    return np.ascontiguousarray(x, dtype=dt)


def out_shape(H, W, k, md, s1, s2, pad):
    oh, ow, oc = _i(), _i(), _i()
    rc = lib().fn2o_correlation_output_shape(H, W, k, md, s1, s2, pad, ctypes.byref(oh), ctypes.byref(ow),
                                             ctypes.byref(oc))
    if rc:
        raise ValueError("correlation geometry rejected (%d)" % rc)
    return oh.value, ow.value, oc.value


# ----------------------------------------------------------------------------------------------- C restatement
def c_correlation(a, b, k, md, s1, s2, pad, f64=False):
    a, b = _c(a), _c(b)
    N, H, W, C = a.shape
    oh, ow, oc = out_shape(H, W, k, md, s1, s2, pad)
    out = np.empty((N, oh, ow, oc), np.float64 if f64 else np.float32)
    fn = lib().fn2o_correlation_forward_f64 if f64 else lib().fn2o_correlation_forward
    rc = fn(a, b, out, N, H, W, C, k, md, s1, s2, pad)
    assert rc == 0, rc
    return out


def c_correlation_grad(g, a, b, k, md, s1, s2, pad, f64=False):
    g, a, b = _c(g), _c(a), _c(b)
    N, H, W, C = a.shape
    dt = np.float64 if f64 else np.float32
    ga, gb = np.empty(a.shape, dt), np.empty(a.shape, dt)
    fn = lib().fn2o_correlation_backward_f64 if f64 else lib().fn2o_correlation_backward
    rc = fn(g, a, b, ga, gb, N, H, W, C, k, md, s1, s2, pad)
    assert rc == 0, rc
    return ga, gb


def c_flow_warp(image, flow):
    image, flow = _c(image), _c(flow)
    N, H, W, C = image.shape
    out = np.empty_like(image)
    assert lib().fn2o_flow_warp_forward(image, flow, out, N, H, W, C) == 0
    return out


def c_flow_warp_grad(image, flow, g):
    image, flow, g = _c(image), _c(flow), _c(g)
    N, H, W, C = image.shape
    gi, gf = np.empty_like(image), np.empty_like(flow)
    assert lib().fn2o_flow_warp_backward(image, flow, g, gi, gf, N, H, W, C) == 0
    return gi, gf


def c_downsample(x, size):
    x = _c(x)
    N, H, W, C = x.shape
    out = np.empty((N, size[0], size[1], C), np.float32)
    rc = lib().fn2o_downsample_forward(x, out, N, H, W, C, size[0], size[1])
    if rc:
        raise ValueError("downsample rejected (%d)" % rc)
    return out


# ------------------------------------------------------------------------------------------- numpy restatement
def _disp(md, s2):
    r = md // s2
    gw = 2 * r + 1
    return r, gw


def np_correlation(a, b, k, md, s1, s2, pad):
    """float64, from the formula of SURVEY.md App. A.1 on explicitly padded arrays."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    N, H, W, C = a.shape
    oh, ow, D = out_shape(H, W, k, md, s1, s2, pad)
    r, gw = _disp(md, s2)
    # pad generously so every read is in range; logical padded coords are offset by `ext`
    ext = md + k + r * s2 + 1
    A = np.zeros((N, H + 2 * pad + 2 * ext, W + 2 * pad + 2 * ext, C))
    B = np.zeros_like(A)
    A[:, ext + pad:ext + pad + H, ext + pad:ext + pad + W] = a
    B[:, ext + pad:ext + pad + H, ext + pad:ext + pad + W] = b
    out = np.zeros((N, oh, ow, D))
    ys = ext + md + s1 * np.arange(oh)
    xs = ext + md + s1 * np.arange(ow)
    for d in range(D):
        dy, dx = (d // gw - r) * s2, (d % gw - r) * s2
        acc = np.zeros((N, oh, ow))
        for j in range(k):
            for i in range(k):
                pa = A[:, ys + j][:, :, xs + i]
                pb = B[:, ys + j + dy][:, :, xs + i + dx]
                acc += (pa * pb).sum(-1)
        out[..., d] = acc / (k * k * C)
    return out


def np_correlation_grad(g, a, b, k, md, s1, s2, pad):
    """float64 adjoint of np_correlation, accumulated into padded gradient arrays (scatter form)."""
    g = np.asarray(g, np.float64)
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    N, H, W, C = a.shape
    oh, ow, D = out_shape(H, W, k, md, s1, s2, pad)
    r, gw = _disp(md, s2)
    ext = md + k + r * s2 + 1
    shape = (N, H + 2 * pad + 2 * ext, W + 2 * pad + 2 * ext, C)
    A, B = np.zeros(shape), np.zeros(shape)
    A[:, ext + pad:ext + pad + H, ext + pad:ext + pad + W] = a
    B[:, ext + pad:ext + pad + H, ext + pad:ext + pad + W] = b
    GA, GB = np.zeros(shape), np.zeros(shape)
    ys = ext + md + s1 * np.arange(oh)
    xs = ext + md + s1 * np.arange(ow)
    iy, ix = np.meshgrid(ys, xs, indexing="ij")
    for d in range(D):
        dy, dx = (d // gw - r) * s2, (d % gw - r) * s2
        gd = g[..., d][..., None] / (k * k * C)
        for j in range(k):
            for i in range(k):
                pa = A[:, iy + j, ix + i]
                pb = B[:, iy + j + dy, ix + i + dx]
                for n in range(N):
                    np.add.at(GA[n], (iy + j, ix + i), gd[n] * pb[n])
                    np.add.at(GB[n], (iy + j + dy, ix + i + dx), gd[n] * pa[n])
    sl = (slice(None), slice(ext + pad, ext + pad + H), slice(ext + pad, ext + pad + W))
    return GA[sl], GB[sl]


def _taps(flow, H, W):
    N = flow.shape[0]
    yy, xx = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    x2 = xx[None].astype(np.float32) + flow[..., 0].astype(np.float32)
    y2 = yy[None].astype(np.float32) + flow[..., 1].astype(np.float32)
    with np.errstate(invalid="ignore"):
        inside = (x2 >= 0) & (y2 >= 0) & (x2 < W) & (y2 < H)
    x2s, y2s = np.where(inside, x2, 0), np.where(inside, y2, 0)
    L, T = x2s.astype(np.int64), y2s.astype(np.int64)
    R, B = np.minimum(L + 1, W - 1), np.minimum(T + 1, H - 1)
    al, be = (x2s - L).astype(np.float64), (y2s - T).astype(np.float64)
    nn = np.broadcast_to(np.arange(N)[:, None, None], L.shape)
    return inside, nn, L, T, R, B, al, be, x2s.astype(np.float64), y2s.astype(np.float64)


def np_flow_warp(image, flow):
    image = np.asarray(image, np.float64)
    N, H, W, C = image.shape
    inside, nn, L, T, R, B, al, be, _, _ = _taps(np.asarray(flow), H, W)
    out = ((1 - al) * (1 - be))[..., None] * image[nn, T, L] + (al * (1 - be))[..., None] * image[nn, T, R] \
        + ((1 - al) * be)[..., None] * image[nn, B, L] + (al * be)[..., None] * image[nn, B, R]
    return np.where(inside[..., None], out, 0.0)


def np_flow_warp_grad(image, flow, g):
    image = np.asarray(image, np.float64)
    g = np.where(np.isfinite(g), g, 0).astype(np.float64)
    N, H, W, C = image.shape
    inside, nn, L, T, R, B, al, be, x2, y2 = _taps(np.asarray(flow), H, W)
    gm = g * inside[..., None]
    gi = np.zeros_like(image)
    np.add.at(gi, (nn, T, L), gm * ((1 - al) * (1 - be))[..., None])
    np.add.at(gi, (nn, T, R), gm * (al * (1 - be))[..., None])
    np.add.at(gi, (nn, B, L), gm * ((1 - al) * be)[..., None])
    np.add.at(gi, (nn, B, R), gm * (al * be)[..., None])
    gy, gx = (B - y2)[..., None], (R - x2)[..., None]
    fx = (gm * (gy * (image[nn, T, R] - image[nn, T, L]) + (1 - gy) * (image[nn, B, R] - image[nn, B, L]))).sum(-1)
    fy = (gm * (gx * (image[nn, B, L] - image[nn, T, L]) + (1 - gx) * (image[nn, B, R] - image[nn, T, R]))).sum(-1)
    return gi, np.stack([fx, fy], -1)


def _round_half_away(x):
    return np.where(x >= 0, np.floor(x + 0.5), np.ceil(x - 0.5)).astype(np.int64)


def np_downsample(x, size):
    """float64 accumulation; index/weight arithmetic in float32 exactly as the kernel does it."""
    x = np.asarray(x, np.float32)
    N, iH, iW, C = x.shape
    oH, oW = size
    f = np.float32
    ws, hs = f(iW - 1) / f(oW - 1), f(iH - 1) / f(oH - 1)
    wr, hr = int(np.ceil(ws)), int(np.ceil(hs))
    srcx = (np.arange(oW, dtype=f) / f(oW - 1)) * f(iW - 1)
    srcy = (np.arange(oH, dtype=f) / f(oH - 1)) * f(iH - 1)
    isx, isy = _round_half_away(srcx), _round_half_away(srcy)
    av = np.zeros((N, oH, oW, C))
    aw = np.zeros((N, oH, oW, C))
    an = np.zeros((N, oH, oW, C))
    for ddy in range(-hr, hr + 1):
        yo = isy + ddy
        oky = (yo >= 0) & (yo < iH)
        wy = np.maximum(f(0), f(1) - np.abs(yo.astype(f) - srcy) / hs)
        for ddx in range(-wr, wr + 1):
            xo = isx + ddx
            okx = (xo >= 0) & (xo < iW)
            wx = np.maximum(f(0), f(1) - np.abs(xo.astype(f) - srcx) / ws)
            w = (wx[None, :] * wy[:, None]).astype(np.float64) * (oky[:, None] & okx[None, :])
            s = x[:, np.clip(yo, 0, iH - 1)][:, :, np.clip(xo, 0, iW - 1)].astype(np.float64)
            isn = np.isnan(s)
            wb = w[None, :, :, None]
            an += np.where(isn, wb, 0)
            av += np.where(isn, 0, s * wb)
            aw += np.where(isn, 0, wb)
    with np.errstate(invalid="ignore", divide="ignore"):
        res = np.where(an / aw > 0.5, np.nan, av / aw)
    return res
