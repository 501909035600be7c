"""Middlebury ``.flo`` optical-flow files, the on-disk format of the reference's ground truth
and outputs: ``src/flowlib.py:85-100`` (read_flow) and ``:127-142`` (write_flow).

Layout: float32 magic 202021.25 ("PIEH"), int32 width, int32 height, then height*width*2 float32, row-major,
(u, v) interleaved -- i.e. one [H, W, 2] NHWC image, exactly what ``flow_warp`` and ``downsample`` take.
"""
import numpy as np

FLO_MAGIC = 202021.25


def read_flo(filename):
    """-> float32 array [H, W, 2].  A wrong magic number raises (the reference prints a message and returns None,
    flowlib.py:93-94)."""
    with open(filename, "rb") as f:
        magic = np.fromfile(f, np.float32, count=1)
        if magic.size != 1 or magic[0] != np.float32(FLO_MAGIC):
            raise ValueError("Magic number incorrect. Invalid .flo file: %s" % filename)
        w = int(np.fromfile(f, np.int32, count=1)[0])
        h = int(np.fromfile(f, np.int32, count=1)[0])
        if w <= 0 or h <= 0:
            raise ValueError("Invalid .flo size %d x %d: %s" % (w, h, filename))
        data = np.fromfile(f, np.float32, count=2 * w * h)
        if data.size != 2 * w * h:
            raise ValueError("Truncated .flo file: %s" % filename)
    return data.reshape(h, w, 2)


def write_flo(flow, filename):
    """flow: array-like [H, W, 2] (torch tensors are taken through .cpu().numpy())."""
    if hasattr(flow, "detach"):
        flow = flow.detach().cpu().numpy()
    flow = np.ascontiguousarray(flow, dtype=np.float32)
    if flow.ndim != 3 or flow.shape[2] != 2:
        raise ValueError("flow must be [H, W, 2]")
    h, w = flow.shape[:2]
    with open(filename, "wb") as f:
        np.array([FLO_MAGIC], np.float32).tofile(f)
        np.array([w], np.int32).tofile(f)
        np.array([h], np.int32).tofile(f)
        flow.tofile(f)
