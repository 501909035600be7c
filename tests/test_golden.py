"""Golden vectors = outputs of the reference's OWN CUDA kernels (tests/golden/make_golden.py ran them on a B200 from
oracle/_ref/libfn2ref.so).  CPU part: the oracle reproduces them.  GPU part: so does the CUDA path."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import oracle as o
from conftest import assert_close

GOLD = os.path.dirname(os.path.abspath(__file__)) + "/golden"
CORR = sorted(glob.glob(GOLD + "/corr_*.npz"))


def test_golden_files_present():
    assert len(CORR) >= 4 and os.path.exists(GOLD + "/warp.npz") and os.path.exists(GOLD + "/downsample.npz")


@pytest.mark.parametrize("path", CORR, ids=[os.path.basename(p) for p in CORR])
def test_oracle_reproduces_reference_correlation(path):
    z = np.load(path)
    k, md, s1, s2, pad = [int(v) for v in z["args"]]
    out = o.c_correlation(z["a"], z["b"], k, md, s1, s2, pad)
    assert_close(out, z["out"], what="forward")
    ga, gb = o.c_correlation_grad(z["g"], z["a"], z["b"], k, md, s1, s2, pad)
    assert_close(ga, z["ga"], what="backprops_a")
    assert_close(gb, z["gb"], what="backprops_b")
    # the double-accumulation restatement agrees too, so fp32 round-off is the only difference
    assert_close(o.c_correlation(z["a"], z["b"], k, md, s1, s2, pad, f64=True), z["out"], what="forward f64")


def test_oracle_reproduces_reference_flow_warp():
    z = np.load(GOLD + "/warp.npz")
    assert_close(o.c_flow_warp(z["image"], z["flow"]), z["out"], what="warp")
    gi, gf = o.c_flow_warp_grad(z["image"], z["flow"], z["g"])
    assert_close(gi, z["gi"], rel=1e-5, abs_=2e-6, what="image_grad")
    assert_close(gf, z["gf"], rel=1e-5, abs_=2e-6, what="flow_grad")


def test_oracle_reproduces_reference_downsample():
    z = np.load(GOLD + "/downsample.npz")
    for key in z.files:
        if key.startswith("out_"):
            size = tuple(int(v) for v in key.split("_")[1:])
            assert_close(o.c_downsample(z["x"], size), z[key], rel=1e-5, abs_=5e-6, what=key)


@pytest.mark.gpu
@pytest.mark.parametrize("path", CORR, ids=[os.path.basename(p) for p in CORR])
def test_cuda_reproduces_reference_correlation(path):
    import flownet2_tf_b200 as fn2
    z = np.load(path)
    k, md, s1, s2, pad = [int(v) for v in z["args"]]
    a, b, g = [torch.from_numpy(z[n]).cuda() for n in ("a", "b", "g")]
    assert_close(fn2.correlation(a, b, k, md, s1, s2, pad).cpu().numpy(), z["out"], what="forward")
    ga, gb = fn2.correlation_grad(g, a, b, k, md, s1, s2, pad)
    assert_close(ga.cpu().numpy(), z["ga"], what="backprops_a")
    assert_close(gb.cpu().numpy(), z["gb"], what="backprops_b")


@pytest.mark.gpu
def test_cuda_reproduces_reference_warp_and_downsample():
    import flownet2_tf_b200 as fn2
    z = np.load(GOLD + "/warp.npz")
    im, fl, g = [torch.from_numpy(z[n]).cuda() for n in ("image", "flow", "g")]
    assert_close(fn2.flow_warp(im, fl).cpu().numpy(), z["out"], what="warp")
    gi, gf = fn2.flow_warp_grad(im, fl, g)
    assert_close(gi.cpu().numpy(), z["gi"], rel=1e-5, abs_=2e-6, what="image_grad")
    assert_close(gf.cpu().numpy(), z["gf"], rel=1e-5, abs_=2e-6, what="flow_grad")
    z = np.load(GOLD + "/downsample.npz")
    x = torch.from_numpy(z["x"]).cuda()
    for key in z.files:
        if key.startswith("out_"):
            size = [int(v) for v in key.split("_")[1:]]
            assert_close(fn2.downsample(x, size).cpu().numpy(), z[key], rel=1e-5, abs_=5e-6, what=key)


@pytest.mark.parametrize("name,flow_min", [("sample0.npz", -108.90625), ("sample1.npz", -37.6875)])
def test_both_oracles_agree_on_the_reference_sample_pair(name, flow_min):
    """tests/golden/sampleK.npz = the reference's data/samples pair K (made by tests/golden/make_fixtures.py): the C
    restatement and the independent float64 numpy restatement give the same warp, gradients and downsampled flow on
    real data (smooth flows, displacements down to -108.9 px, pixel values 0..255)."""
    z = np.load(os.path.join(GOLD, name))
    assert z["image"].shape == (384, 512, 3) and z["flow"].shape == (384, 512, 2)
    assert float(z["flow"].min()) == flow_min
    image = z["image"].astype(np.float32)[None, 100:228, 200:392]
    flow = np.ascontiguousarray(z["flow"][None, 100:228, 200:392])
    assert_close(o.c_flow_warp(image, flow), o.np_flow_warp(image, flow), rel=1e-5, abs_=1e-4, what="warp")
    g = np.random.default_rng(0).standard_normal(image.shape).astype(np.float32)
    ci, cf = o.c_flow_warp_grad(image, flow, g)
    ni, nf = o.np_flow_warp_grad(image, flow, g)
    assert_close(ci, ni, rel=1e-5, abs_=2e-5, what="image_grad")
    assert_close(cf, nf, rel=1e-5, abs_=2e-3, what="flow_grad")
    for size in [(6, 8), (24, 32)]:
        assert_close(o.c_downsample(flow, size), o.np_downsample(flow, size), rel=1e-5, abs_=5e-6, what=str(size))
