"""Weight import / export for the caller graphs: the reference's two on-disk weight formats
into (and out of) the ``state_dict`` of ``flownet2_tf_b200.graphs``.

* **Caffe-npy** -- ``scripts/caffe/convert_caffe_weights_to_npy.py:488-497`` pickles ONE dict
  ``{tf_scope + '/weights': HWIO array, tf_scope + '/biases': [O]}`` with ``np.save``; Caffe's ``(out, in, h, w)``
  kernels were transposed ``(2, 3, 1, 0)`` into TensorFlow's ``(h, w, in, out)`` there (``:491-495``).  The same
  permutation applied to a Caffe deconvolution ``(in, out, h, w)`` gives slim.conv2d_transpose's ``(h, w, out, in)``.
* **TF checkpoint (V2 "tensor bundle")** -- what ``tf.train.Saver`` writes and ``slim.assign_from_checkpoint`` reads
  (``src/net.py:1424-1448``): ``<prefix>.index`` is an SSTable (LevelDB table format) mapping variable names to
  BundleEntryProto records (dtype, shape, shard, offset, size, masked crc32c), ``<prefix>.data-00000-of-0000N`` holds
  the raw little-endian tensor bytes.  Read and written here without TensorFlow.

Variable names follow the reference's variable scopes (``flownet_c.py:17``, ``flownet_cs.py:16``, ``flownet2.py:20``...):
``FlowNet2/FlowNetCSS/FlowNetCS/FlowNetC/conv1/weights``.  ``net.py:1426-1432`` restores a sub-net's checkpoint into a
stacked net by cutting the names at ``new_scope + '/'``; ``rescope`` does the same to a name dict.

Going back to PyTorch: conv ``(h, w, in, out) -> (out, in, h, w)`` and transposed conv ``(h, w, out, in) ->
(in, out, h, w)`` are both ``transpose(3, 2, 0, 1)``.
"""
import os
import pickle
import struct

import numpy as np
import torch

# module class name -> TF variable scope it opens; attribute names of sub-nets are resolved through the class
_SCOPES = {"FlowNetC": "FlowNetC", "FlowNetS": "FlowNetS", "FlowNetSD": "FlowNetSD", "FlowNetCS": "FlowNetCS",
           "FlowNetCSS": "FlowNetCSS", "FlowNet2": "FlowNet2"}


def tf_variable_names(model):
    """{state_dict key: TF variable name} for a graph of ``flownet2_tf_b200.graphs``.

    ``refine`` (the shared decoder) opens no scope; ``conv``/``deconv`` leaf modules are the slim layers, whose
    variables are ``<scope>/weights`` and ``<scope>/biases``."""
    names = {}

    def walk(mod, prefix, scope):
        for attr, child in mod.named_children():
            cls = type(child).__name__
            if cls in _SCOPES:
                walk(child, prefix + attr + ".", scope + _SCOPES[cls] + "/")
            elif cls == "Refinement":
                walk(child, prefix + attr + ".", scope)
            elif cls in ("Conv", "Deconv"):
                inner = "conv" if cls == "Conv" else "deconv"
                layer = getattr(child, inner)
                names[prefix + attr + "." + inner + ".weight"] = scope + attr + "/weights"
                if layer.bias is not None:
                    names[prefix + attr + "." + inner + ".bias"] = scope + attr + "/biases"
            else:
                walk(child, prefix + attr + ".", scope)

    root = _SCOPES.get(type(model).__name__)
    if root is None:
        raise TypeError("not one of the caller graphs: %s" % type(model).__name__)
    walk(model, "", root + "/")
    missing = set(model.state_dict()) - set(names)
    if missing:
        raise RuntimeError("state_dict entries without a TF name: %s" % sorted(missing))
    return names


def to_torch_layout(array):
    """HWIO (conv) / HWOI (transposed conv) -> PyTorch's OIHW / IOHW; 1-D arrays (biases) pass through."""
    array = np.asarray(array)
    return np.ascontiguousarray(array.transpose(3, 2, 0, 1)) if array.ndim == 4 else np.ascontiguousarray(array)


def to_tf_layout(array):
    """The inverse of ``to_torch_layout`` (``convert_caffe_weights_to_npy.py:494``: ``transpose((2, 3, 1, 0))``)."""
    array = np.asarray(array)
    return np.ascontiguousarray(array.transpose(2, 3, 1, 0)) if array.ndim == 4 else np.ascontiguousarray(array)


def rescope(variables, scope, new_scope):
    """``net.py:1424-1432``: keep the variables under ``scope`` and name them by what follows ``new_scope + '/'`` --
    how the reference loads e.g. a FlowNetC checkpoint (names ``FlowNetC/...``) into ``FlowNetCS/FlowNetC/...``:
    ``rescope(names_of_stacked_net, 'FlowNetCS/FlowNetC', 'FlowNetCS')``.  Works on any dict keyed by variable name."""
    out = {}
    for name, value in variables.items():
        if name == scope or name.startswith(scope + "/"):
            out[name.split(new_scope + "/")[1]] = value
    return out


def state_dict_from_variables(model, variables, strict=True, prefix_map=None):
    """Builds a ``state_dict`` for ``model`` from ``{TF variable name: array in TF layout}``.

    ``prefix_map``: optional ``{scope in the model: scope in the file}``, e.g. ``{'FlowNetCS/FlowNetC': 'FlowNetC'}`` to
    take the C tower of a stacked net from a stand-alone FlowNetC file (the renaming of ``net.py:1424-1432``).
    ``strict=False`` skips variables the file does not have (``optimistic_restore_vars``, ``net.py:37-65``)."""
    names = tf_variable_names(model)
    ref = model.state_dict()
    out, missing = {}, []
    for key, tfname in names.items():
        src = tfname
        for mine, theirs in (prefix_map or {}).items():
            if tfname == mine or tfname.startswith(mine + "/"):
                src = theirs + tfname[len(mine):]
                break
        if src not in variables:
            missing.append(src)
            continue
        arr = to_torch_layout(variables[src])
        if tuple(arr.shape) != tuple(ref[key].shape):
            raise ValueError("%s: file has %s (torch layout %s), the graph wants %s" % (
                src, tuple(np.shape(variables[src])), tuple(arr.shape), tuple(ref[key].shape)))
        out[key] = torch.from_numpy(arr.astype(np.float32, copy=False)).clone()
    if missing and strict:
        raise KeyError("variables not in the file: %s" % sorted(missing)[:8])
    return out


def variables_from_state_dict(model):
    """{TF variable name: float32 array in TF layout} of the model's current weights."""
    names = tf_variable_names(model)
    sd = model.state_dict()
    return {tfname: to_tf_layout(sd[key].detach().cpu().numpy()) for key, tfname in names.items()}


# ------------------------------------------------------------------------------------------------ Caffe-npy
_NPY_PICKLE_GLOBALS = {   # all that a pickled {name: ndarray} dict needs, under numpy 1.x and 2.x module paths
    ("numpy.core.multiarray", "_reconstruct"), ("numpy._core.multiarray", "_reconstruct"),
    ("numpy.core.multiarray", "scalar"), ("numpy._core.multiarray", "scalar"),
    ("numpy", "ndarray"), ("numpy", "dtype"), ("_codecs", "encode"), ("collections", "OrderedDict"),
}


class _ArraysOnlyUnpickler(pickle.Unpickler):
    """pickle.Unpickler that resolves nothing but the constructors of numpy arrays, dtypes and scalars."""

    def find_class(self, module, name):
        if (module, name) in _NPY_PICKLE_GLOBALS:
            return super().find_class(module, name)
        raise pickle.UnpicklingError(
            "weights file asks for %s.%s: not a plain {name: array} dict (refused; pass trusted=True only for a file "
            "whose origin you trust)" % (module, name))


def load_npy(path, trusted=False):
    """The pickled dict written by ``convert_caffe_weights_to_npy.py:497`` (``np.save(path, dict)``).

    That format IS a pickle: ``np.load(allow_pickle=True)`` on a downloaded weights file would run whatever code the
    file names.  By default the file is therefore read with an unpickler that only admits numpy arrays, dtypes and
    scalars inside builtin containers, which is all the reference's converter writes; anything else is refused.
    ``trusted=True`` hands the file to ``np.load(allow_pickle=True)`` as the reference does -- arbitrary code execution
    if the file is hostile.  ``convert_npy_to_npz`` rewrites a file once into a pickle-free ``.npz``."""
    if trusted:
        obj = np.load(path, allow_pickle=True)
    else:
        with open(path, "rb") as f:
            version = np.lib.format.read_magic(f)
            if version == (1, 0):
                _, _, dtype = np.lib.format.read_array_header_1_0(f)
            else:
                _, _, dtype = np.lib.format.read_array_header_2_0(f)
            if not dtype.hasobject:
                raise ValueError("%s does not hold a {name: array} dict" % path)
            try:
                obj = _ArraysOnlyUnpickler(f).load()
            except UnicodeError:      # written by Python 2 (the Caffe tool chain): 8-bit strings
                f.seek(0)
                np.lib.format.read_magic(f)
                (np.lib.format.read_array_header_1_0 if version == (1, 0) else np.lib.format.read_array_header_2_0)(f)
                obj = _ArraysOnlyUnpickler(f, encoding="latin1").load()
    if isinstance(obj, np.ndarray) and obj.dtype == object and obj.shape == ():
        obj = obj.item()
    if not isinstance(obj, dict):
        raise ValueError("%s does not hold a {name: array} dict" % path)
    return {str(k): np.asarray(v) for k, v in obj.items()}


def convert_npy_to_npz(path, out_path, trusted=False):
    """Rewrites the pickled dict as a plain ``.npz`` archive (no pickle inside); ``load_npz`` reads it back."""
    np.savez(out_path, **load_npy(path, trusted=trusted))


def load_npz(path):
    with np.load(path, allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def save_npy(path, variables):
    with open(path, "wb") as f:   # a file object: np.save would otherwise append ".npy"
        np.save(f, dict(variables), allow_pickle=True)


def load_npy_into(model, path, strict=True, prefix_map=None, trusted=False):
    """``path``: the reference's pickled ``.npy`` dict (read with the arrays-only unpickler unless ``trusted``; see
    ``load_npy``) or a ``.npz`` written by ``convert_npy_to_npz``."""
    variables = load_npz(path) if str(path).endswith(".npz") else load_npy(path, trusted=trusted)
    model.load_state_dict(state_dict_from_variables(model, variables, strict, prefix_map), strict=strict)
    return model


# --------------------------------------------------------------------------------------- TF checkpoint V2 bundle
_TABLE_MAGIC = 0xdb4775248b80fb57
_DT = {1: np.dtype("<f4"), 2: np.dtype("<f8"), 3: np.dtype("<i4"), 9: np.dtype("<i8"), 10: np.dtype("bool")}
_DT_INV = {np.dtype("float32"): 1, np.dtype("float64"): 2, np.dtype("int32"): 3, np.dtype("int64"): 9}


def _crc_table():
    poly, tab = 0x82F63B78, []           # CRC-32C (Castagnoli), reflected
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ poly if c & 1 else c >> 1
        tab.append(c)
    return np.array(tab, dtype=np.uint32)


_CRC_TAB = _crc_table()


def _crc_shift_tables(nbytes):
    """Four 256-entry tables T such that the raw crc register r after feeding ``nbytes`` zero bytes is
    T[0][r & 255] ^ T[1][(r >> 8) & 255] ^ ... (the register update is linear over GF(2))."""
    # images of the 32 unit vectors under one zero byte, then square-and-multiply on the 32x32 bit matrix
    def apply(mat, v):
        out, i = 0, 0
        while v:
            if v & 1:
                out ^= mat[i]
            v >>= 1
            i += 1
        return out

    one = [int((_CRC_TAB[(1 << b) & 0xff] if b < 8 else 0) ^ ((1 << b) >> 8)) for b in range(32)]
    result, power, n = [1 << b for b in range(32)], one, nbytes
    while n:
        if n & 1:
            result = [apply(power, col) for col in result]
        power = [apply(power, col) for col in power]
        n >>= 1
    tabs = np.zeros((4, 256), dtype=np.uint32)
    for k in range(4):
        for v in range(256):
            tabs[k, v] = apply(result, v << (8 * k))
    return tabs


def crc32c(data):
    """CRC-32C of a bytes-like object.  Large buffers are cut into 4 KiB pieces whose registers advance together as
    numpy vectors; the pieces are then joined with the shift-by-4096-zero-bytes operator."""
    buf = np.frombuffer(memoryview(data).cast("B"), dtype=np.uint8)
    n = buf.size
    piece = 4096
    crc = np.uint32(0xffffffff)
    whole = n // piece
    if whole >= 8:
        blocks = buf[:whole * piece].reshape(whole, piece)
        reg = np.zeros(whole, dtype=np.uint32)          # raw registers, started at 0 (linearity: the init is added below)
        for j in range(piece):
            reg = _CRC_TAB[(reg ^ blocks[:, j]) & 0xff] ^ (reg >> 8)
        tabs = _crc_shift_tables(piece)
        r = int(crc)
        for k in range(whole):
            r = int(tabs[0][r & 255] ^ tabs[1][(r >> 8) & 255] ^ tabs[2][(r >> 16) & 255] ^ tabs[3][r >> 24]) ^ int(reg[k])
        crc = np.uint32(r)
        tail = buf[whole * piece:]
    else:
        tail = buf
    r = int(crc)
    tab = _CRC_TAB
    for b in tail.tolist():
        r = int(tab[(r ^ b) & 0xff]) ^ (r >> 8)
    return r ^ 0xffffffff


def _mask(crc):
    return ((((crc >> 15) | (crc << 17)) & 0xffffffff) + 0xa282ead8) & 0xffffffff


def _varint(buf, pos):
    out, shift = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7f) << shift
        if not b & 0x80:
            return out, pos
        shift += 7


def _put_varint(v):
    out = bytearray()
    while True:
        b = v & 0x7f
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _proto_fields(buf):
    """Minimal protobuf wire decoder: yields (field number, wire type, value); length-delimited values as bytes."""
    pos = 0
    while pos < len(buf):
        key, pos = _varint(buf, pos)
        field, wt = key >> 3, key & 7
        if wt == 0:
            val, pos = _varint(buf, pos)
        elif wt == 1:
            val = struct.unpack_from("<Q", buf, pos)[0]
            pos += 8
        elif wt == 2:
            ln, pos = _varint(buf, pos)
            val = bytes(buf[pos:pos + ln])
            pos += ln
        elif wt == 5:
            val = struct.unpack_from("<I", buf, pos)[0]
            pos += 4
        else:
            raise ValueError("unsupported protobuf wire type %d" % wt)
        yield field, wt, val


def _snappy_uncompress(buf):
    n, pos = _varint(buf, 0)
    out = bytearray()
    while pos < len(buf):
        tag = buf[pos]
        pos += 1
        kind = tag & 3
        if kind == 0:
            ln = (tag >> 2) + 1
            if ln > 60:
                nb = ln - 60
                ln = int.from_bytes(buf[pos:pos + nb], "little") + 1
                pos += nb
            out += buf[pos:pos + ln]
            pos += ln
            continue
        if kind == 1:
            ln = ((tag >> 2) & 7) + 4
            off = ((tag >> 5) << 8) | buf[pos]
            pos += 1
        elif kind == 2:
            ln = (tag >> 2) + 1
            off = int.from_bytes(buf[pos:pos + 2], "little")
            pos += 2
        else:
            ln = (tag >> 2) + 1
            off = int.from_bytes(buf[pos:pos + 4], "little")
            pos += 4
        for _ in range(ln):
            out.append(out[-off])
    if len(out) != n:
        raise ValueError("corrupt snappy block")
    return bytes(out)


def _read_block(raw, offset, size):
    body, kind = raw[offset:offset + size], raw[offset + size]
    stored = struct.unpack_from("<I", raw, offset + size + 1)[0]
    if stored != _mask(crc32c(raw[offset:offset + size + 1])):
        raise ValueError("checkpoint index: block checksum mismatch")
    if kind == 1:
        body = _snappy_uncompress(body)
    elif kind != 0:
        raise ValueError("checkpoint index: unknown block compression %d" % kind)
    nrestarts = struct.unpack_from("<I", body, len(body) - 4)[0]
    end = len(body) - 4 - 4 * nrestarts
    pos, key, out = 0, b"", []
    while pos < end:
        shared, pos = _varint(body, pos)
        non_shared, pos = _varint(body, pos)
        vlen, pos = _varint(body, pos)
        key = key[:shared] + bytes(body[pos:pos + non_shared])
        pos += non_shared
        out.append((key, bytes(body[pos:pos + vlen])))
        pos += vlen
    return out


def read_tf_checkpoint_index(prefix):
    """-> (num_shards, {name: dict(dtype, shape, shard, offset, size, crc32c)}) from ``<prefix>.index``."""
    raw = open(prefix + ".index", "rb").read()
    if len(raw) < 48 or struct.unpack_from("<Q", raw, len(raw) - 8)[0] != _TABLE_MAGIC:
        raise ValueError("%s.index is not a TensorFlow V2 checkpoint index (bad table magic)" % prefix)
    footer = raw[len(raw) - 48:]
    _, p = _varint(footer, 0)          # metaindex handle (unused)
    _, p = _varint(footer, p)
    ioff, p = _varint(footer, p)
    isize, p = _varint(footer, p)
    entries, shards = {}, 1
    for _, handle in _read_block(raw, ioff, isize):
        boff, q = _varint(handle, 0)
        bsize, q = _varint(handle, q)
        for key, val in _read_block(raw, boff, bsize):
            if key == b"":                           # BundleHeaderProto: num_shards = 1, endianness = 2
                for f, _, v in _proto_fields(val):
                    if f == 1:
                        shards = v
                    if f == 2 and v != 0:
                        raise ValueError("big-endian checkpoints are not supported")
                continue
            e = {"dtype": 0, "shape": (), "shard": 0, "offset": 0, "size": 0, "crc32c": None, "sliced": False}
            for f, _, v in _proto_fields(val):
                if f == 1:
                    e["dtype"] = v
                elif f == 2:
                    dims = []
                    for f2, _, v2 in _proto_fields(v):
                        if f2 == 2:
                            size = 0
                            for f3, _, v3 in _proto_fields(v2):
                                if f3 == 1:
                                    size = v3
                            dims.append(size)
                    e["shape"] = tuple(dims)
                elif f == 3:
                    e["shard"] = v
                elif f == 4:
                    e["offset"] = v
                elif f == 5:
                    e["size"] = v
                elif f == 6:
                    e["crc32c"] = v
                elif f == 7:
                    e["sliced"] = True
            entries[key.decode("utf-8")] = e
    return shards, entries


def load_tf_checkpoint(prefix, verify=True, only=None):
    """{variable name: array} of a V2 checkpoint (``<prefix>.index`` + ``<prefix>.data-*``), names as saved.
    ``only``: optional predicate on the name (the optimiser's slot variables ``.../Adam`` are rarely wanted).
    ``verify`` checks every tensor's crc32c."""
    shards, entries = read_tf_checkpoint_index(prefix)
    files, out = {}, {}
    for name, e in sorted(entries.items()):
        if only is not None and not only(name):
            continue
        if e["sliced"]:
            raise ValueError("%s: partitioned variables are not supported" % name)
        if e["dtype"] not in _DT:
            continue                                  # strings etc.: not weights
        if e["shard"] not in files:
            files[e["shard"]] = np.memmap("%s.data-%05d-of-%05d" % (prefix, e["shard"], shards), dtype=np.uint8, mode="r")
        raw = files[e["shard"]][e["offset"]:e["offset"] + e["size"]]
        if verify and e["crc32c"] is not None and _mask(crc32c(raw)) != e["crc32c"]:
            raise ValueError("%s: tensor checksum mismatch" % name)
        out[name] = np.frombuffer(raw.tobytes(), dtype=_DT[e["dtype"]]).reshape(e["shape"]).copy()
    return out


def _block(entries):
    """One table block, every entry a restart point (no prefix sharing) + trailer (no compression, masked crc)."""
    body, restarts = bytearray(), []
    for key, val in entries:
        restarts.append(len(body))
        body += _put_varint(0) + _put_varint(len(key)) + _put_varint(len(val)) + key + val
    for r in restarts or [0]:
        body += struct.pack("<I", r)
    body += struct.pack("<I", max(1, len(restarts)))
    body += b"\x00"
    return bytes(body) + struct.pack("<I", _mask(crc32c(bytes(body))))


def _field(num, wt, payload):
    return _put_varint((num << 3) | wt) + payload


def save_tf_checkpoint(prefix, variables, global_step=None):
    """Writes ``{name: array}`` as a one-shard V2 checkpoint in the layout ``tf.train.Saver`` /
    ``slim.assign_from_checkpoint`` read (and ``load_tf_checkpoint`` reads back).  TensorFlow-side compatibility is
    UNTESTED: there is no TensorFlow in this image and no TF-written checkpoint under the reference tree, so nothing
    here has been read by TensorFlow, and the reader has only seen files built to the published format (this writer,
    and the hand-built prefix-compressed and snappy blocks of tests/test_weights.py).  ``global_step`` adds the int64 scalar the
    reference's training loop expects (``net.py:1443`` parses the step from the file name as well)."""
    items = {k: np.ascontiguousarray(v) for k, v in variables.items()}
    if global_step is not None:
        items["global_step"] = np.array(int(global_step), dtype=np.int64)
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    header = _field(1, 0, _put_varint(1)) + _field(3, 2, _put_varint(2) + _field(1, 0, _put_varint(1)))  # shards, version
    entries, offset = [(b"", header)], 0
    with open(prefix + ".data-00000-of-00001", "wb") as data:
        for name in sorted(items):
            arr = items[name]
            if arr.dtype not in _DT_INV:
                raise TypeError("%s: dtype %s is not supported" % (name, arr.dtype))
            raw = arr.astype(arr.dtype.newbyteorder("<"), copy=False).tobytes()
            shape = b"".join(_field(2, 2, (lambda d: _put_varint(len(d)) + d)(_field(1, 0, _put_varint(int(s)))))
                             for s in arr.shape)
            entry = _field(1, 0, _put_varint(_DT_INV[arr.dtype])) + _field(2, 2, _put_varint(len(shape)) + shape)
            if offset:
                entry += _field(4, 0, _put_varint(offset))
            entry += _field(5, 0, _put_varint(len(raw))) + _field(6, 5, struct.pack("<I", _mask(crc32c(raw))))
            entries.append((name.encode("utf-8"), entry))
            data.write(raw)
            offset += len(raw)
    # table: data blocks of <= 64 entries, an empty metaindex block, the index block, the 48-byte footer
    out, index = bytearray(), []
    for i in range(0, len(entries), 64):
        chunk = entries[i:i + 64]
        blk = _block(chunk)
        index.append((chunk[-1][0], _put_varint(len(out)) + _put_varint(len(blk) - 5)))
        out += blk
    meta = _block([])
    meta_handle = _put_varint(len(out)) + _put_varint(len(meta) - 5)
    out += meta
    idx = _block(index)
    idx_handle = _put_varint(len(out)) + _put_varint(len(idx) - 5)
    out += idx
    footer = meta_handle + idx_handle
    out += footer + b"\x00" * (40 - len(footer)) + struct.pack("<Q", _TABLE_MAGIC)
    with open(prefix + ".index", "wb") as f:
        f.write(out)


def load_tf_checkpoint_into(model, prefix, strict=True, prefix_map=None):
    """Restores ``model`` from a V2 checkpoint: variables are taken by the names of ``tf_variable_names`` (optimizer
    slots and ``global_step`` are ignored), HWIO -> OIHW."""
    variables = load_tf_checkpoint(prefix, only=lambda n: n.endswith("/weights") or n.endswith("/biases"))
    model.load_state_dict(state_dict_from_variables(model, variables, strict, prefix_map), strict=strict)
    return model
