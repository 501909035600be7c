"""Packs the reference's own sample pairs into tests/golden/sample0.npz and sample1.npz (inputs only; ~0.5 MB each,
compressed); for pair K:

    image  uint8  [384,512,3]   /root/reference/data/samples/Kimg1.ppm   (the image FlowNet2 warps, flownet_cs.py:24)
    image0 uint8  [384,512,3]   /root/reference/data/samples/Kimg0.ppm   (the reference image of the brightness error)
    flow   float32 [384,512,2]  /root/reference/data/samples/Kflow.flo   (ground truth; pair 0: flows down to -108.9 px)

Run where /root/reference exists:  python tests/golden/make_fixtures.py
The GPU tests (tests/test_gpu_fixtures.py) run the kernels on these real inputs against the oracle and, when built,
against the reference's own kernels; the CPU test (tests/test_golden.py) checks the oracle on them.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
SAMPLES = "/root/reference/data/samples"


def read_ppm(path):
    raw = open(path, "rb").read()
    assert raw[:2] == b"P6", path
    fields, pos = [], 2
    while len(fields) < 3:            # width, height, maxval: whitespace-separated ASCII
        while raw[pos:pos + 1].isspace():
            pos += 1
        end = pos
        while not raw[end:end + 1].isspace():
            end += 1
        fields.append(int(raw[pos:end]))
        pos = end
    pos += 1                          # the single whitespace byte after maxval
    w, h, mx = fields
    assert mx == 255
    return np.frombuffer(raw, np.uint8, count=w * h * 3, offset=pos).reshape(h, w, 3).copy()


def main():
    from flownet2_tf_b200.flo import read_flo
    for k in (0, 1):
        out = os.path.join(ROOT, "tests", "golden", "sample%d.npz" % k)
        np.savez_compressed(out, image=read_ppm(os.path.join(SAMPLES, "%dimg1.ppm" % k)),
                            image0=read_ppm(os.path.join(SAMPLES, "%dimg0.ppm" % k)),
                            flow=read_flo(os.path.join(SAMPLES, "%dflow.flo" % k)))
        print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
