"""GPU parity on the reference's REAL sample pairs (tests/golden/sampleK.npz = data/samples/Kimg1.ppm, Kimg0.ppm,
Kflow.flo for K = 0 and 1, packed by tests/golden/make_fixtures.py; every test runs on both): flow_warp forward / backward and all five loss scales of
downsample on the ground-truth flow (SURVEY.md section 8 c/d), against the oracle and, when built, against the
reference's own kernels.  Real flows are smooth with large displacements (down to -108.9 px): neighbouring pixels share
taps, many taps leave the staged window -- the opposite regime of the synthetic iid flows of the other tests."""
import os

import numpy as np
import pytest
import torch

from conftest import assert_close, grad_like

pytestmark = pytest.mark.gpu

import flownet2_tf_b200 as fn2  # noqa: E402
from oracle import oracle as o  # noqa: E402
from oracle import ref  # noqa: E402

LOSS_SIZES = [(6, 8), (12, 16), (24, 32), (48, 64), (96, 128)]   # flownet_c.py:135-160
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module", params=["sample0.npz", "sample1.npz"])
def sample(request):
    z = np.load(os.path.join(GOLD, request.param))
    image = torch.from_numpy(z["image"].astype(np.float32))[None].contiguous()     # [1,384,512,3], 0..255 as the
    image0 = torch.from_numpy(z["image0"].astype(np.float32))[None].contiguous()   # reference feeds it (net.py:301)
    flow = torch.from_numpy(z["flow"])[None].contiguous()
    return image, image0, flow


def test_flow_warp_forward_and_error_on_the_sample_pair(sample):
    image, image0, flow = sample
    want = o.c_flow_warp(image.numpy(), flow.numpy())
    got = fn2.flow_warp(image.cuda(), flow.cuda())
    assert_close(got.cpu().numpy(), want, what="warp of 0img1 by 0flow vs oracle")
    if ref.available():
        r = ref.flow_warp(image.cuda(), flow.cuda())
        torch.cuda.synchronize()
        assert_close(got.cpu().numpy(), r.cpu().numpy(), what="vs the reference kernel")
    warped, err = fn2.flow_warp_error(image.cuda(), flow.cuda(), image0.cuda())
    assert torch.equal(warped, got)
    assert_close(err.cpu().numpy()[..., 0], np.sqrt(((image0.numpy() - want) ** 2).sum(-1)), rel=2e-5, abs_=1e-4,
                 what="brightness error")


def test_flow_warp_backward_on_the_sample_pair(sample):
    image, _, flow = sample
    g = grad_like(image.shape)
    gi, gf = fn2.flow_warp_grad(image.cuda(), flow.cuda(), g.cuda())
    wi, wf = o.c_flow_warp_grad(image.numpy(), flow.numpy(), g.numpy())
    # image gradient: fp32 atomics in an arbitrary order, here as in the reference (flow_warp_grad.cu.cc:43-53)
    assert_close(gi.cpu().numpy(), wi, rel=1e-5, abs_=2e-5, what="image_grad vs oracle")
    # flow gradient: differences of pixel values up to 255 times the incoming gradient
    assert_close(gf.cpu().numpy(), wf, rel=1e-5, abs_=2e-3, what="flow_grad vs oracle")
    if ref.available():
        r_gi, r_gf = ref.flow_warp_grad(image.cuda(), flow.cuda(), g.cuda())
        torch.cuda.synchronize()
        assert_close(gi.cpu().numpy(), r_gi.cpu().numpy(), rel=1e-5, abs_=2e-5, what="image_grad vs reference kernel")
        assert_close(gf.cpu().numpy(), r_gf.cpu().numpy(), rel=1e-5, abs_=2e-3, what="flow_grad vs reference kernel")
    # 16 copies with the pair rolled differently: the batched launch gives every image its own answer
    shifts = [(7 * i) % 384 for i in range(4)]
    ims = torch.cat([image.roll(s, 1) for s in shifts]).cuda()
    fls = torch.cat([flow.roll(s, 1) for s in shifts]).cuda()
    gs = torch.cat([g.roll(s, 1) for s in shifts]).cuda()
    bi, bf = fn2.flow_warp_grad(ims, fls, gs)
    for k in range(len(shifts)):
        si, sf = fn2.flow_warp_grad(ims[k:k + 1].contiguous(), fls[k:k + 1].contiguous(), gs[k:k + 1].contiguous())
        assert_close(bi[k:k + 1].cpu().numpy(), si.cpu().numpy(), rel=1e-5, abs_=2e-5, what="batch slice %d" % k)
        assert torch.equal(bf[k:k + 1], sf)


@pytest.mark.parametrize("scale", [1.0, 0.05])
def test_downsample_of_the_ground_truth_flow_all_loss_scales(sample, scale):
    """Real flows reach 109 px and the coarsest scale averages 153 x 147 samples: the reference's own sequential
    fp32 sum is 8.5e-5 away from the exact value there (C restatement vs the float64 restatement, which agree with the
    reference kernel / with exact arithmetic respectively), so the value check is against the float64 oracle, with the
    absolute term scaled by the data (5e-7 * max|x|), and the fp32-order comparisons get 2e-6 * max|x|."""
    _, _, flow = sample
    fs = (flow * scale).contiguous()       # the loss downsamples flow * 0.05 (flownet_c.py:128)
    mag = float(fs.abs().max())
    single = [fn2.downsample(fs.cuda(), list(s)) for s in LOSS_SIZES]
    multi = fn2.downsample_multi(fs.cuda(), LOSS_SIZES)
    for size, a, b in zip(LOSS_SIZES, single, multi):
        exact = o.np_downsample(fs.numpy(), size)
        assert_close(a.cpu().numpy(), exact, rel=1e-5, abs_=5e-7 * mag, what="downsample %s vs float64" % (size,))
        assert_close(b.cpu().numpy(), exact, rel=1e-5, abs_=5e-7 * mag, what="downsample_multi %s vs float64" % (size,))
        want = o.c_downsample(fs.numpy(), size)
        assert_close(b.cpu().numpy(), want, rel=1e-5, abs_=2e-6 * mag, what="downsample_multi %s vs fp32 order" % (size,))
        if ref.available():
            r = ref.downsample(fs.cuda(), size)
            torch.cuda.synchronize()
            assert_close(want, r.cpu().numpy(), rel=1e-5, abs_=2e-7 * mag, what="oracle vs reference kernel %s" % (size,))
            assert_close(b.cpu().numpy(), r.cpu().numpy(), rel=1e-5, abs_=2e-6 * mag, what="vs reference kernel %s" % (size,))


def test_multiscale_epe_on_the_ground_truth_flow(sample):
    _, _, flow = sample
    g = torch.Generator().manual_seed(3)
    preds = [torch.randn(1, h, w, 2, generator=g) for h, w in LOSS_SIZES]
    epe = fn2.multiscale_epe(flow.cuda(), [p.cuda() for p in preds], 0.05)
    fs = (flow * 0.05).numpy()
    for k, (size, p) in enumerate(zip(LOSS_SIZES, preds)):
        d = o.np_downsample(fs, size).astype(np.float64)
        want = np.sqrt(((p.numpy() - d) ** 2).sum(-1)).sum() / 1.0     # utils.py:219-224
        assert abs(float(epe[k]) - want) <= 1e-5 * abs(want) + 1e-5, (size, float(epe[k]), want)
