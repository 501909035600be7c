"""Weight import / export: Caffe-npy dicts and TF V2 checkpoints <-> state_dict, CPU only."""
import os
import re
import struct

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


# ---- files this package's writer did NOT produce -------------------------------------------------------------
# The writer above emits one restart point per entry and no compression; index files written by TensorFlow use
# LevelDB's block builder (keys prefix-compressed against the previous key, a restart point every 16 entries) and may
# snappy-compress blocks.  No TensorFlow-written checkpoint exists under the reference tree (checkpoints/download.sh
# needs the network), so these tests build such blocks by hand, from the published table / snappy formats, with
# encoders that live here and share no code with weights.py's writer.
def _varint_t(v):
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7f) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _leveldb_block(entries, restart_interval=16):
    body, restarts, prev, count = bytearray(), [], b"", 0
    for key, val in entries:
        shared = 0
        if count % restart_interval == 0:
            restarts.append(len(body))
        else:
            while shared < min(len(prev), len(key)) and prev[shared] == key[shared]:
                shared += 1
        body += _varint_t(shared) + _varint_t(len(key) - shared) + _varint_t(len(val)) + key[shared:] + val
        prev, count = key, count + 1
    for r in restarts or [0]:
        body += struct.pack("<I", r)
    body += struct.pack("<I", max(1, len(restarts)))
    return bytes(body)


def _snappy_t(data):
    """A small greedy snappy encoder: literals (short and long form), copies with 1-, 2- and 4-byte offsets."""
    out, i, lit, table = bytearray(_varint_t(len(data))), 0, bytearray(), {}

    def flush():
        nonlocal lit
        while lit:
            chunk, lit = lit[:70000], lit[70000:]
            n = len(chunk) - 1
            if n < 60:
                out.append(n << 2)
            else:
                nb = (n.bit_length() + 7) // 8
                out.append((59 + nb) << 2)
                out.extend(n.to_bytes(nb, "little"))
            out.extend(chunk)

    while i < len(data):
        key = data[i:i + 4]
        j = table.get(key) if len(key) == 4 else None
        table[key] = i
        if j is None:
            lit.append(data[i])
            i += 1
            continue
        ln = 4
        while i + ln < len(data) and ln < 64 and data[j + ln] == data[i + ln]:   # j + ln may run into i: overlapping copy
            ln += 1
        off = i - j
        flush()
        if ln <= 11 and off < 2048:
            out.append(1 | ((ln - 4) << 2) | ((off >> 8) << 5))
            out.append(off & 0xff)
        elif off < 65536:
            out.append(2 | ((ln - 1) << 2))
            out.extend(off.to_bytes(2, "little"))
        else:
            out.append(3 | ((ln - 1) << 2))
            out.extend(off.to_bytes(4, "little"))
        i += ln
    flush()
    return bytes(out)


def _sealed(body, compress):
    raw = (_snappy_t(body) if compress else body) + (b"\x01" if compress else b"\x00")
    return raw + struct.pack("<I", weights._mask(weights.crc32c(raw)))


def _pb(num, wt, payload):
    return _varint_t((num << 3) | wt) + payload


def _hand_built_checkpoint(prefix, tensors, compress):
    """<prefix>.index / .data-00000-of-00001 for {name: float32 array}: data blocks of 40 entries with prefix-shared
    keys (restart interval 16), optionally snappy-compressed, then metaindex, index block and footer."""
    header = _pb(1, 0, _varint_t(1)) + _pb(3, 2, _varint_t(2) + _pb(1, 0, _varint_t(1)))
    entries, offset = [(b"", header)], 0
    with open(prefix + ".data-00000-of-00001", "wb") as f:
        for name in sorted(tensors):
            raw = tensors[name].astype("<f4").tobytes()
            dims = b"".join(_pb(2, 2, _varint_t(len(d)) + d) for d in (_pb(1, 0, _varint_t(n)) for n in tensors[name].shape))
            e = _pb(1, 0, _varint_t(1)) + _pb(2, 2, _varint_t(len(dims)) + dims)
            e += _pb(4, 0, _varint_t(offset)) + _pb(5, 0, _varint_t(len(raw)))
            e += _pb(6, 5, struct.pack("<I", weights._mask(weights.crc32c(raw))))
            entries.append((name.encode(), e))
            f.write(raw)
            offset += len(raw)
    out, index = bytearray(), []
    for i in range(0, len(entries), 40):
        chunk = entries[i:i + 40]
        body = _leveldb_block(chunk)
        blk = _sealed(body, compress)
        index.append((chunk[-1][0], _varint_t(len(out)) + _varint_t(len(blk) - 5)))
        out += blk
    meta = _sealed(_leveldb_block([]), False)
    handles = _varint_t(len(out)) + _varint_t(len(meta) - 5)
    out += meta
    idx = _sealed(_leveldb_block(index, restart_interval=1), False)
    handles += _varint_t(len(out)) + _varint_t(len(idx) - 5)
    out += idx
    out += handles + b"\x00" * (40 - len(handles)) + struct.pack("<Q", 0xdb4775248b80fb57)
    with open(prefix + ".index", "wb") as f:
        f.write(out)


@pytest.mark.parametrize("compress", [False, True])
def test_reader_takes_prefix_compressed_and_snappy_blocks(tmp_path, compress):
    rng = np.random.default_rng(3)
    names = ["FlowNet2/FlowNetCSS/FlowNetCS/FlowNetC/conv%d%s/%s" % (i, suffix, kind)
             for i in range(1, 7) for suffix in ("", "_1") for kind in ("weights", "biases", "weights/Adam", "weights/Adam_1")]
    names += ["FlowNet2/FlowNetSD/deconv%d/weights" % i for i in range(2, 6)] + ["global_step_f"]
    tensors = {n: rng.standard_normal((3, 2, 1 + i % 5)).astype(np.float32) for i, n in enumerate(names)}
    prefix = str(tmp_path / ("hand%d.ckpt-7" % compress))
    _hand_built_checkpoint(prefix, tensors, compress)
    shards, entries = weights.read_tf_checkpoint_index(prefix)
    assert shards == 1 and sorted(entries) == sorted(names)        # 53 keys: two data blocks, restarts every 16 entries
    back = weights.load_tf_checkpoint(prefix)
    for n in names:
        assert np.array_equal(back[n], tensors[n]), n


def test_snappy_reader_on_each_element_kind():
    rng = np.random.default_rng(5)
    cases = [
        b"", b"a", b"abcd" * 3, b"x" * 500,                                    # run-length: copies overlapping their output
        bytes(rng.integers(0, 256, 300, dtype=np.uint8)),                       # one long-form literal
        (b"FlowNetC/conv3_1/weights" + bytes(rng.integers(0, 256, 3000, dtype=np.uint8))) * 3,   # 2-byte offsets
        b"tensorbundle" + bytes(rng.integers(0, 256, 70000, dtype=np.uint8)) + b"tensorbundle-again",   # 4-byte offset
    ]
    for data in cases:
        comp = _snappy_t(data)
        assert weights._snappy_uncompress(comp) == data
    kinds = set()
    mixed = cases[-1] + cases[-2] + cases[3] + b"0123456789!01234567?"        # ends with an 8-byte match 11 bytes back
    comp, pos = _snappy_t(mixed), 0
    assert weights._snappy_uncompress(comp) == mixed
    _, pos = weights._varint(comp, 0)
    while pos < len(comp):                                                      # the streams above really hold all four kinds
        tag = comp[pos]
        kinds.add(tag & 3)
        if tag & 3 == 0:
            ln = (tag >> 2) + 1
            if ln > 60:
                nb = ln - 60
                ln = int.from_bytes(comp[pos + 1:pos + 1 + nb], "little") + 1
                pos += nb
            pos += 1 + ln
        else:
            pos += {1: 2, 2: 3, 3: 5}[tag & 3]
    assert kinds == {0, 1, 2, 3}
    with pytest.raises(ValueError):
        weights._snappy_uncompress(_varint_t(9) + _snappy_t(b"abcdefgh")[1:])   # declared length does not match


def test_npy_loader_refuses_pickles_that_are_not_plain_arrays(tmp_path):
    import pickle
    good = str(tmp_path / "w.npy")
    variables = {"conv1/weights": np.arange(24, dtype=np.float32).reshape(2, 3, 4), "conv1/biases": np.ones(4, np.float32)}
    weights.save_npy(good, variables)
    back = weights.load_npy(good)                                   # default: arrays-only unpickler
    assert sorted(back) == sorted(variables) and np.array_equal(back["conv1/weights"], variables["conv1/weights"])
    assert np.array_equal(weights.load_npy(good, trusted=True)["conv1/biases"], variables["conv1/biases"])

    class Evil:
        def __reduce__(self):
            return (os.system, ("echo pwned > %s" % (tmp_path / "pwned"),))

    bad = str(tmp_path / "bad.npy")
    with open(bad, "wb") as f:
        np.save(f, np.array({"conv1/weights": Evil()}, dtype=object), allow_pickle=True)
    with pytest.raises(pickle.UnpicklingError, match="refused"):
        weights.load_npy(bad)
    assert not (tmp_path / "pwned").exists()
    npz = str(tmp_path / "w.npz")
    weights.convert_npy_to_npz(good, npz)
    assert np.array_equal(weights.load_npz(npz)["conv1/weights"], variables["conv1/weights"])
