"""The caller graphs (flownet2_tf_b200.graphs) on CPU with the oracle standing in for the
kernels: wiring, shapes, the loss of flownet_c.py:127-166, gradients through the ops, and the .flo format (f-4)."""
import os

import numpy as np
import pytest
import torch

from conftest import oracle_ops

import flownet2_tf_b200 as fn2
from flownet2_tf_b200 import graphs as G


def _pair(n=1, h=64, w=64, seed=0):
    g = torch.Generator().manual_seed(seed)
    return {"input_a": torch.rand(n, h, w, 3, generator=g), "input_b": torch.rand(n, h, w, 3, generator=g)}


def test_flownet_c_shapes_and_reference_layer_sizes():
    torch.manual_seed(0)
    net = G.FlowNetC(ops=oracle_ops())
    # layer inventory of flownet_c.py:30-108
    assert net.conv1.conv.weight.shape == (64, 3, 7, 7) and net.conv1.conv.stride == (2, 2)
    assert net.conv3.conv.weight.shape == (256, 128, 5, 5)
    assert net.conv_redir.conv.weight.shape == (32, 256, 1, 1)
    assert net.conv3_1.conv.weight.shape == (256, 473, 3, 3)          # 32 + 441 (flownet_c.py:44-48)
    assert net.refine.deconv5.deconv.bias is None                      # biases_initializer=None (:58)
    assert net.refine.predict_flow5.conv.weight.shape == (2, 512 + 512 + 2, 3, 3)
    assert net.refine.predict_flow2.conv.weight.shape == (2, 128 + 64 + 2, 3, 3)
    out = net(_pair(2, 64, 128))
    assert {k: tuple(v.shape) for k, v in out.items()} == {
        "predict_flow6": (2, 1, 2, 2), "predict_flow5": (2, 2, 4, 2), "predict_flow4": (2, 4, 8, 2),
        "predict_flow3": (2, 8, 16, 2), "predict_flow2": (2, 16, 32, 2), "flow": (2, 64, 128, 2)}
    # flow = predict_flow2 * 20 resized with align_corners=True (:112-116): the corners are exact samples
    pf2, flow = out["predict_flow2"], out["flow"]
    torch.testing.assert_close(flow[:, 0, 0], pf2[:, 0, 0] * 20.0)
    torch.testing.assert_close(flow[:, -1, -1], pf2[:, -1, -1] * 20.0)


def test_correlation_sits_where_the_reference_calls_it():
    calls = []
    ops = oracle_ops()
    inner = ops.correlation

    def spy(a, b, *args):
        calls.append((tuple(a.shape), tuple(b.shape), args))
        return inner(a, b, *args)
    ops.correlation = spy
    torch.manual_seed(1)
    G.FlowNetC(ops=ops)(_pair(1, 64, 64))
    assert calls == [((1, 8, 8, 256), (1, 8, 8, 256), (1, 20, 1, 2, 20))]   # conv3 of both towers, flownet_c.py:40


def test_stacks_wire_warp_and_brightness_error():
    torch.manual_seed(2)
    ops = oracle_ops()
    warps = []
    inner = ops.flow_warp
    ops.flow_warp = lambda image, flow: (warps.append((image, flow)), inner(image, flow))[1]
    inputs = _pair(1, 64, 64, seed=3)
    cs = G.FlowNetCS(ops=ops)
    assert cs.net_s.conv1.conv.weight.shape == (64, 12, 7, 7)           # flownet_s.py:18-24
    out = cs(inputs)
    assert tuple(out["flow"].shape) == (1, 64, 64, 2) and len(warps) == 1
    assert warps[0][0] is inputs["input_b"] or torch.equal(warps[0][0], inputs["input_b"])   # warps image B (:21)
    # the frozen FlowNetC receives no gradient (trainable=False, flownet_cs.py:18), FlowNetS does
    out["flow"].sum().backward()
    assert all(p.grad is None for p in cs.net_c.parameters())
    assert cs.net_s.conv1.conv.weight.grad is not None
    warps.clear()
    net2 = G.FlowNet2(ops=ops)
    assert net2.fuse_conv0.conv.weight.shape == (64, 11, 3, 3)          # flownet2.py:41-47
    assert net2.fuse_deconv1.deconv.bias is not None                     # the fusion net's deconvs keep biases
    out2 = net2(inputs)
    assert tuple(out2["predict_flow0"].shape) == (1, 64, 64, 2) and tuple(out2["flow"].shape) == (1, 64, 64, 2)
    assert len(warps) == 4                                               # C->S, CS->S, and the two fusion warps


def test_default_ops_are_the_cuda_ops_and_refuse_cpu_tensors():
    ops = G.default_ops("tf32")
    assert ops.flow_warp is fn2.flow_warp and ops.downsample_multi is fn2.downsample_multi
    with pytest.raises(ValueError, match="CUDA"):
        ops.correlation(torch.zeros(1, 4, 4, 32), torch.zeros(1, 4, 4, 32), 1, 20, 1, 2, 20)


def test_flownet_sd_scale_and_interconv():
    torch.manual_seed(4)
    net = G.FlowNetSD()
    assert net.refine.interconv2.conv.weight.shape == (64, 128 + 64 + 2, 3, 3)
    out = net(_pair(1, 64, 64))
    torch.testing.assert_close(out["flow"][:, 0, 0], out["predict_flow2"][:, 0, 0] * 0.05)   # flownet_sd.py:106


def test_loss_matches_a_direct_restatement():
    torch.manual_seed(5)
    ops = oracle_ops()
    g = torch.Generator().manual_seed(6)
    gt = torch.randn(2, 64, 64, 2, generator=g) * 4
    preds = {"predict_flow%d" % k: torch.randn(2, 64 >> s, 64 >> s, 2, generator=g)
             for k, s in zip((6, 5, 4, 3, 2), (6, 5, 4, 3, 2))}
    preds["predict_flow6"] = torch.randn(2, 2, 2, 2, generator=g)       # downsample needs >= 2 per side
    preds["predict_flow5"] = torch.randn(2, 2, 2, 2, generator=g)
    total, parts = G.multiscale_loss(gt, preds, ops)
    want = 0.0
    for wgt, k in zip((0.32, 0.08, 0.02, 0.01, 0.005), (6, 5, 4, 3, 2)):
        p = preds["predict_flow%d" % k]
        d = ops.downsample(gt * 0.05, [p.shape[1], p.shape[2]])
        want += wgt * float(torch.sqrt(((p - d) ** 2).sum(3)).sum() / 2)
    want /= 5       # tf.losses.compute_weighted_loss, default reduction SUM_BY_NONZERO_WEIGHTS (flownet_c.py:163)
    assert abs(float(total) - want) < 1e-4 * max(1.0, abs(want))
    assert len(parts) == 5


def test_hfem_losses_match_a_numpy_restatement():
    """average_endpoint_error_hfem (utils.py:227-339) in its three modes and the multi-scale loss built on it
    (flownet_s_interp.py:160-245), against float64 numpy written from the formulas."""
    ops = oracle_ops()
    g = torch.Generator().manual_seed(16)
    gt = torch.randn(2, 64, 64, 2, generator=g) * 4
    edges = torch.rand(2, 64, 64, 1, generator=g)
    sides = (2, 2, 4, 8, 16)
    preds = {"predict_flow%d" % k: torch.randn(2, n, n, 2, generator=g) for k, n in zip((6, 5, 4, 3, 2), sides)}
    lam, perc = 2.0, 50
    for mode in ("edges", "hard", ""):
        total, parts = G.multiscale_loss_hfem(gt, preds, ops, add_hfem=mode, lambda_w=lam, perc_hfem=perc, edges=edges)
        want = []
        for k in (6, 5, 4, 3, 2):
            p = preds["predict_flow%d" % k].numpy().astype(np.float64)
            size = [p.shape[1], p.shape[2]]
            d = ops.downsample(gt * 0.05, size).numpy().astype(np.float64)
            epe = np.sqrt(((p - d) ** 2).sum(3, keepdims=True))
            if mode == "edges":
                e = ops.downsample(edges, size).numpy().astype(np.float64)
                want.append((epe + lam * epe * e).sum() / 2)
            elif mode == "hard":
                flat = np.sort(epe.reshape(-1))[::-1]
                kk = int(round(perc / 100 * flat.size))
                want.append((1 + lam) * flat[:kk].sum() / 2 * flat.size / kk)
            else:
                want.append(epe.sum() / 2)
        np.testing.assert_allclose([float(x) for x in parts], want, rtol=2e-5)
        wsum = sum(w * l for w, l in zip((0.32, 0.08, 0.02, 0.01, 0.005), want)) / 5
        assert abs(float(total) - wsum) < 1e-4 * max(1.0, abs(wsum))
    # 'edges' without an edge map is the plain loss (utils.py:335-337)
    a, _ = G.multiscale_loss_hfem(gt, preds, ops, add_hfem="edges", edges=None)
    b, _ = G.multiscale_loss(gt, preds, ops)
    assert float(a) == float(b)


def test_weighted_loss_is_pinned_to_a_hand_computed_compute_weighted_loss():
    """tf.losses.compute_weighted_loss(losses, weights) = sum(losses * weights) / count(weights != 0): for losses
    10, 20, 30, 40, 50 and the FlowNet weights that is (3.2 + 1.6 + 0.6 + 0.4 + 0.25) / 5 = 1.21, not 6.05."""
    losses = [torch.tensor(v) for v in (10.0, 20.0, 30.0, 40.0, 50.0)]
    assert abs(float(G.weighted_loss(losses, G.LOSS_WEIGHTS)) - 1.21) < 1e-6
    assert abs(float(G.weighted_loss(losses, (0.5, 0.0, 0.0, 0.0, 0.25))) - (5.0 + 12.5) / 2) < 1e-6   # zero weights do not count


def test_training_step_sends_gradients_through_the_correlation():
    torch.manual_seed(7)
    ops = oracle_ops()
    net = G.FlowNetC(ops=ops)
    inputs = _pair(1, 128, 128, seed=8)
    gt = torch.randn(1, 128, 128, 2, generator=torch.Generator().manual_seed(9))
    out = net(inputs)
    loss, _ = G.multiscale_loss(gt, out, ops)
    loss = loss + G.l2_regularization(net, 4e-4)
    loss.backward()
    for name in ("conv1", "conv3", "conv_redir", "conv3_1"):
        gr = getattr(net, name).conv.weight.grad
        assert gr is not None and torch.isfinite(gr).all() and float(gr.abs().sum()) > 0, name


def test_flo_round_trip_and_header(tmp_path):
    flow = np.random.RandomState(0).randn(5, 7, 2).astype(np.float32)
    path = str(tmp_path / "x.flo")
    fn2.write_flo(flow, path)
    raw = open(path, "rb").read()
    assert raw[:4] == b"PIEH" and len(raw) == 12 + 5 * 7 * 2 * 4          # flowlib.py:127-142
    assert np.frombuffer(raw[4:12], np.int32).tolist() == [7, 5]          # width, then height
    np.testing.assert_array_equal(fn2.read_flo(path), flow)
    fn2.write_flo(torch.from_numpy(flow), path)
    np.testing.assert_array_equal(fn2.read_flo(path), flow)
    open(path, "wb").write(b"XXXX" + raw[4:])
    with pytest.raises(ValueError, match="Magic"):
        fn2.read_flo(path)
    with pytest.raises(ValueError):
        fn2.write_flo(np.zeros((3, 3, 3), np.float32), path)


@pytest.mark.skipif(not os.path.exists("/root/reference/data/samples/0flow.flo"), reason="reference fixtures absent")
def test_flo_reads_the_reference_sample():
    f = fn2.read_flo("/root/reference/data/samples/0flow.flo")
    assert f.shape == (384, 512, 2) and np.isfinite(f).all()
