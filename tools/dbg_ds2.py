import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import flownet2_tf_b200 as fn2
from oracle import oracle as o
size = (96, 128)
np.set_printoptions(linewidth=250, precision=7)
for (y0, x0) in [(383, 386), (200, 300), (5, 7)]:
    x = torch.zeros(1, 384, 512, 2)
    x[0, y0, x0, 0] = 1.0
    got = fn2.downsample(x.cuda(), list(size)).cpu().numpy()[0, :, :, 0]
    want = o.c_downsample(x.numpy(), size)[0, :, :, 0]
    nz = np.argwhere((got != 0) | (want != 0))
    print("impulse", (y0, x0))
    for (dy, dx) in nz:
        print("   ", (dy, dx), got[dy, dx], want[dy, dx], got[dy, dx] - want[dy, dx])
