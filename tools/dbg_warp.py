import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import flownet2_tf_b200 as fn2
C = int(sys.argv[1]) if len(sys.argv) > 1 else 3
img = torch.rand(1, 64, 64, C, device="cuda")
flo = torch.randn(1, 64, 64, 2, device="cuda") * 4
g = torch.randn(1, 64, 64, C, device="cuda")
out = fn2.flow_warp(img, flo); torch.cuda.synchronize(); print("fwd ok")
gi, gf = fn2.flow_warp_grad(img, flo, g); torch.cuda.synchronize(); print("bwd ok", float(gi.abs().sum()), float(gf.abs().sum()))
