"""Weight import / export: Caffe-npy dicts and TF V2 checkpoints <-> state_dict, CPU only."""
import os
import re

import numpy as np
import pytest
import torch

import flownet2_tf_b200  # noqa: F401
from flownet2_tf_b200 import graphs, weights
from conftest import oracle_ops

REF_CONVERT = "/root/reference/scripts/caffe/convert_caffe_weights_to_npy.py"


def test_crc32c_known_answers_and_vector_path():
    assert weights.crc32c(b"123456789") == 0xE3069283          # the CRC-32C check value
    assert weights.crc32c(b"") == 0
    assert weights.crc32c(bytes(32)) == 0x8A9136AA              # RFC 3720 B.4: 32 zero bytes
    data = np.random.default_rng(0).integers(0, 256, 100_003, dtype=np.uint8).tobytes()   # > 8 pieces + ragged tail
    r = 0xffffffff
    for b in data:                                              # byte-at-a-time definition
        r = int(weights._CRC_TAB[(r ^ b) & 0xff]) ^ (r >> 8)
    assert weights.crc32c(data) == r ^ 0xffffffff


def test_variable_names_follow_the_reference_scopes():
    c = graphs.FlowNetC(ops=oracle_ops())
    names = weights.tf_variable_names(c)
    assert names["conv1.conv.weight"] == "FlowNetC/conv1/weights"
    assert names["conv_redir.conv.bias"] == "FlowNetC/conv_redir/biases"
    assert names["refine.deconv5.deconv.weight"] == "FlowNetC/deconv5/weights"
    assert "refine.deconv5.deconv.bias" not in names                       # biases_initializer=None (flownet_c.py:57)
    f2 = graphs.FlowNet2(ops=oracle_ops())
    n2 = set(weights.tf_variable_names(f2).values())
    assert "FlowNet2/FlowNetCSS/FlowNetCS/FlowNetC/conv3_1/weights" in n2
    assert "FlowNet2/FlowNetCSS/FlowNetCS/FlowNetS/conv1/weights" in n2
    assert "FlowNet2/FlowNetCSS/FlowNetS/upsample_flow6to5/weights" in n2
    assert "FlowNet2/FlowNetSD/interconv5/biases" in n2
    assert "FlowNet2/fuse_deconv1/biases" in n2                            # the fusion net's deconvs do have biases
    assert len(n2) == len(f2.state_dict())


@pytest.mark.skipif(not os.path.exists(REF_CONVERT), reason="reference checkout absent")
def test_names_cover_the_reference_conversion_table():
    """Every TF scope the reference's Caffe conversion writes (convert_caffe_weights_to_npy.py PARAMS tables) is a
    variable of the corresponding graph here."""
    src = open(REF_CONVERT).read()
    scopes = set(re.findall(r":\s*'((?:FlowNet[A-Za-z0-9]*/)+[A-Za-z0-9_]+)'", src))
    assert len(scopes) > 100
    models = {"FlowNetC": graphs.FlowNetC, "FlowNetS": graphs.FlowNetS, "FlowNetSD": graphs.FlowNetSD,
              "FlowNetCS": graphs.FlowNetCS, "FlowNetCSS": graphs.FlowNetCSS, "FlowNet2": graphs.FlowNet2}
    have = {}
    for root in sorted({s.split("/")[0] for s in scopes}):
        m = models[root](ops=oracle_ops()) if root not in ("FlowNetS", "FlowNetSD") else models[root]()
        have[root] = {n.rsplit("/", 1)[0] for n in weights.tf_variable_names(m).values()}
    missing = [s for s in scopes if s not in have[s.split("/")[0]]]
    assert not missing, missing[:10]


def test_layout_is_the_inverse_of_the_conversion_transpose():
    w = np.arange(2 * 3 * 5 * 7, dtype=np.float32).reshape(2, 3, 5, 7)    # Caffe / torch (out, in, h, w)
    tf_layout = w.transpose((2, 3, 1, 0))                                 # convert_caffe_weights_to_npy.py:494
    assert np.array_equal(weights.to_tf_layout(w), tf_layout)
    assert np.array_equal(weights.to_torch_layout(tf_layout), w)
    assert weights.to_torch_layout(np.ones(4)).shape == (4,)


def test_npy_round_trip_and_forward_equality(tmp_path):
    torch.manual_seed(0)
    a = graphs.FlowNetS()
    b = graphs.FlowNetS()
    path = str(tmp_path / "flownet_s.npy")
    weights.save_npy(path, weights.variables_from_state_dict(a))
    loaded = weights.load_npy(path)
    assert loaded["FlowNetS/conv1/weights"].shape == (7, 7, 6, 64)        # HWIO
    assert loaded["FlowNetS/deconv5/weights"].shape == (4, 4, 512, 1024)  # (h, w, out, in)
    weights.load_npy_into(b, path)
    for k, v in a.state_dict().items():
        assert torch.equal(v, b.state_dict()[k]), k
    x = {"input_a": torch.rand(1, 64, 64, 3), "input_b": torch.rand(1, 64, 64, 3)}
    assert torch.equal(a(x)["flow"], b(x)["flow"])


def test_tf_checkpoint_round_trip(tmp_path):
    torch.manual_seed(1)
    a = graphs.FlowNetC(ops=oracle_ops())
    prefix = str(tmp_path / "ckpt" / "model.ckpt-1200000")
    variables = weights.variables_from_state_dict(a)
    variables["FlowNetC/conv1/weights/Adam"] = np.zeros((7, 7, 3, 64), np.float32)   # an optimiser slot, to be ignored
    weights.save_tf_checkpoint(prefix, variables, global_step=1200000)
    shards, entries = weights.read_tf_checkpoint_index(prefix)
    assert shards == 1 and entries["global_step"]["dtype"] == 9 and entries["global_step"]["shape"] == ()
    assert entries["FlowNetC/conv3_1/weights"]["shape"] == (3, 3, 473, 256)
    back = weights.load_tf_checkpoint(prefix)
    assert int(back["global_step"]) == 1200000
    assert np.array_equal(back["FlowNetC/conv6_1/weights"], variables["FlowNetC/conv6_1/weights"])
    b = graphs.FlowNetC(ops=oracle_ops())
    weights.load_tf_checkpoint_into(b, prefix)
    for k, v in a.state_dict().items():
        assert torch.equal(v, b.state_dict()[k]), k
    # a flipped byte in the data file is caught by the per-tensor crc32c
    data = prefix + ".data-00000-of-00001"
    raw = bytearray(open(data, "rb").read())
    raw[len(raw) // 2] ^= 0x40
    open(data, "wb").write(raw)
    with pytest.raises(ValueError, match="checksum"):
        weights.load_tf_checkpoint(prefix)


def test_sub_net_checkpoint_into_a_stacked_net(tmp_path):
    """net.py:1424-1432: a stand-alone FlowNetC checkpoint restores the C tower of FlowNetCS by scope renaming."""
    torch.manual_seed(2)
    c = graphs.FlowNetC(ops=oracle_ops())
    path = str(tmp_path / "c.npy")
    weights.save_npy(path, weights.variables_from_state_dict(c))
    cs = graphs.FlowNetCS(ops=oracle_ops())
    sd = weights.state_dict_from_variables(cs, weights.load_npy(path), strict=False,
                                           prefix_map={"FlowNetCS/FlowNetC": "FlowNetC"})
    assert set(sd) == {"net_c." + k for k in c.state_dict()}
    cs.load_state_dict(sd, strict=False)
    for k, v in c.state_dict().items():
        assert torch.equal(v, cs.net_c.state_dict()[k])
    # the same renaming expressed as the reference does it
    names = {v: k for k, v in weights.tf_variable_names(cs).items()}
    renamed = weights.rescope(names, "FlowNetCS/FlowNetC", "FlowNetCS")
    assert renamed["FlowNetC/conv1/weights"] == "net_c.conv1.conv.weight"
    with pytest.raises(KeyError):
        weights.state_dict_from_variables(cs, weights.load_npy(path), strict=True,
                                          prefix_map={"FlowNetCS/FlowNetC": "FlowNetC"})
