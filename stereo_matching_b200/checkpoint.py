"""Dependency-free reader/writer for the TensorFlow "TensorBundle" checkpoints
the reference ships (`new_checkpoint/ckpt-1000.{index,data-00000-of-00001}`).

The reference restores its weights with
`tf.train.Checkpoint(step, optimizer, net=model).restore(manager.latest_checkpoint)`
(/root/reference/inference.py:86-91, train.py:78-87).  TensorFlow is not a
dependency of this library, so the on-disk format is parsed here directly:

* `<prefix>.index` is a LevelDB-style SSTable: a 48-byte footer (two varint
  block handles + magic 0xdb4775248b80fb57), an index block whose values are
  handles of data blocks, and data blocks of prefix-compressed
  (shared, non_shared, value_len, key_delta, value) entries followed by a
  restart array.  Every block is trailed by 1 compression byte + 4 CRC bytes.
* key ""  -> BundleHeaderProto, every other key -> BundleEntryProto
  {1:dtype 2:shape{2:dim{1:size}} 3:shard_id 4:offset 5:size 6:crc32c}.
* `<prefix>.data-00000-of-00001` holds the raw little-endian tensors.

`write_bundle` emits the same format (single data block, no compression).
`tf.train.Checkpoint.restore` is object-based: it walks the `TrackableObjectGraph`
stored in the string tensor `_CHECKPOINTABLE_OBJECT_GRAPH` and looks tensors up by
the `checkpoint_key` of each node, so a bundle without that tensor would be
restored by NAME (`conv2d/kernel` ...), find nothing and leave the weights at
their initial values.  `object_graph` builds the graph of
`tf.train.Checkpoint(step, optimizer, net)` (train.py:80) for a trunk of N
conv/BN blocks - for base_model_ws_9 byte for byte the one inside the shipped
ckpt-1000 (tests/test_checkpoint.py) - and `write_bundle` stores `bytes` values as
scalar DT_STRING tensors the way TensorFlow does.
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass
from typing import Dict, List, Tuple

import numpy as np

_MAGIC = 0xDB4775248B80FB57

# tensorflow/core/framework/types.proto enum values used by the shipped bundle
_DT_NUMPY = {1: np.dtype("<f4"), 2: np.dtype("<f8"), 3: np.dtype("<i4"),
             9: np.dtype("<i8"), 10: np.dtype("bool")}
_DT_STRING = 7
_NUMPY_DT = {np.dtype("<f4"): 1, np.dtype("<f8"): 2, np.dtype("<i4"): 3,
             np.dtype("<i8"): 9, np.dtype("bool"): 10}


class CheckpointError(ValueError):
    pass


def _varint(buf: bytes, pos: int) -> Tuple[int, int]:
    out = 0
    shift = 0
    while True:
        if pos >= len(buf):
            raise CheckpointError("truncated varint")
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7
        if shift > 63:
            raise CheckpointError("varint too long")


def _put_varint(v: int) -> bytes:
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _block_entries(block: bytes) -> List[Tuple[bytes, bytes]]:
    """Decode one SSTable block (without its 5-byte trailer)."""
    if len(block) < 4:
        raise CheckpointError("block too small")
    n_restarts = struct.unpack_from("<I", block, len(block) - 4)[0]
    end = len(block) - 4 - 4 * n_restarts
    if end < 0:
        raise CheckpointError("bad restart array")
    pos = 0
    key = b""
    out = []
    while pos < end:
        shared, pos = _varint(block, pos)
        non_shared, pos = _varint(block, pos)
        vlen, pos = _varint(block, pos)
        if shared > len(key):
            raise CheckpointError("bad prefix compression")
        key = key[:shared] + block[pos:pos + non_shared]
        pos += non_shared
        out.append((key, block[pos:pos + vlen]))
        pos += vlen
    return out


def _proto_fields(buf: bytes):
    """Yield (field_number, wire_type, value) for a protobuf message."""
    pos = 0
    while pos < len(buf):
        tag, pos = _varint(buf, pos)
        field, wt = tag >> 3, tag & 7
        if wt == 0:
            v, pos = _varint(buf, pos)
        elif wt == 1:
            v = buf[pos:pos + 8]
            pos += 8
        elif wt == 2:
            n, pos = _varint(buf, pos)
            v = buf[pos:pos + n]
            pos += n
        elif wt == 5:
            v = buf[pos:pos + 4]
            pos += 4
        else:
            raise CheckpointError(f"unsupported wire type {wt}")
        yield field, wt, v


@dataclass
class BundleEntry:
    dtype: int
    shape: Tuple[int, ...]
    shard_id: int
    offset: int
    size: int
    crc32c: int


def _parse_entry(buf: bytes) -> BundleEntry:
    dtype = 0
    shape: List[int] = []
    shard = offset = size = crc = 0
    for f, wt, v in _proto_fields(buf):
        if f == 1:
            dtype = v
        elif f == 2:
            for f2, _, v2 in _proto_fields(v):
                if f2 == 2:  # dim
                    dim = 0
                    for f3, _, v3 in _proto_fields(v2):
                        if f3 == 1:
                            dim = v3
                    shape.append(dim)
        elif f == 3:
            shard = v
        elif f == 4:
            offset = v
        elif f == 5:
            size = v
        elif f == 6:
            crc = struct.unpack("<I", v)[0]
    return BundleEntry(dtype, tuple(shape), shard, offset, size, crc)


# ---- crc32c (Castagnoli), as TF stores it masked ---------------------------
_CRC_TABLE = None


def _crc_table():
    global _CRC_TABLE
    if _CRC_TABLE is None:
        tab = []
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
            tab.append(c)
        _CRC_TABLE = tab
    return _CRC_TABLE


def crc32c(data: bytes) -> int:
    tab = _crc_table()
    c = 0xFFFFFFFF
    for b in data:
        c = tab[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc32c(data: bytes) -> int:
    c = crc32c(data)
    return ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


# ---- reading ---------------------------------------------------------------
def read_index(index_path: str) -> Dict[str, BundleEntry]:
    """Parse `<prefix>.index` into {key: BundleEntry} (header key '' dropped)."""
    with open(index_path, "rb") as f:
        buf = f.read()
    if len(buf) < 48:
        raise CheckpointError(f"{index_path}: too small for an SSTable footer")
    footer = buf[-48:]
    if struct.unpack("<Q", footer[40:])[0] != _MAGIC:
        raise CheckpointError(f"{index_path}: bad SSTable magic")
    pos = 0
    _, pos = _varint(footer, pos)      # metaindex offset
    _, pos = _varint(footer, pos)      # metaindex size
    idx_off, pos = _varint(footer, pos)
    idx_size, pos = _varint(footer, pos)
    entries: Dict[str, BundleEntry] = {}
    for _, handle in _block_entries(buf[idx_off:idx_off + idx_size]):
        off, p = _varint(handle, 0)
        size, p = _varint(handle, p)
        if buf[off + size] != 0:
            raise CheckpointError("compressed SSTable blocks are not supported")
        for key, val in _block_entries(buf[off:off + size]):
            if key == b"":
                continue
            entries[key.decode("utf-8")] = _parse_entry(val)
    return entries


def latest_checkpoint(model_dir: str) -> str | None:
    """Mirror of `tf.train.CheckpointManager.latest_checkpoint`
    (/root/reference/inference.py:89-91): read `<model_dir>/checkpoint`."""
    state = os.path.join(model_dir, "checkpoint")
    if not os.path.exists(state):
        return None
    with open(state) as f:
        for line in f:
            if line.startswith("model_checkpoint_path:"):
                name = line.split(":", 1)[1].strip().strip('"')
                return name if os.path.isabs(name) else os.path.join(model_dir, name)
    return None


def _string_payload(value: bytes) -> Tuple[bytes, int]:
    """On-disk bytes of a scalar DT_STRING tensor and the entry's crc32c (tensor_bundle.cc WriteStringTensor):
    varint length, masked crc32c of the length as a little-endian uint32, the bytes; the entry checksum runs
    over the FIXED-width length, the length checksum and the bytes (checked against ckpt-1000's object graph)."""
    n = struct.pack("<I", len(value))
    len_crc = struct.pack("<I", masked_crc32c(n))
    return _put_varint(len(value)) + len_crc + value, masked_crc32c(n + len_crc + value)


def _read_string(raw: bytes, entry: "BundleEntry", key: str, verify_crc: bool) -> bytes:
    n, pos = _varint(raw, 0)
    value = raw[pos + 4:]
    if len(value) != n:
        raise CheckpointError(f"{key}: string tensor length mismatch")
    if verify_crc:
        payload, crc = _string_payload(value)
        if payload != raw or crc != entry.crc32c:
            raise CheckpointError(f"{key}: crc32c mismatch")
    return value


def read_bundle(prefix: str, verify_crc: bool = False,
                keys_filter=None, strings: bool = False) -> Dict[str, np.ndarray]:
    """Read every numeric tensor of a bundle; scalar string tensors are returned as `bytes` when `strings` is
    set and skipped otherwise."""
    entries = read_index(prefix + ".index")
    data_path = prefix + ".data-00000-of-00001"
    with open(data_path, "rb") as f:
        data = f.read()
    out: Dict[str, np.ndarray] = {}
    for key, e in entries.items():
        if keys_filter is not None and not keys_filter(key):
            continue
        if e.dtype == _DT_STRING and strings and not e.shape:
            out[key] = _read_string(data[e.offset:e.offset + e.size], e, key, verify_crc)
            continue
        if e.dtype == _DT_STRING or e.dtype not in _DT_NUMPY:
            continue
        if e.shard_id != 0:
            raise CheckpointError("multi-shard bundles are not supported")
        raw = data[e.offset:e.offset + e.size]
        if len(raw) != e.size:
            raise CheckpointError(f"{key}: data file truncated")
        if verify_crc and masked_crc32c(raw) != e.crc32c:
            raise CheckpointError(f"{key}: crc32c mismatch")
        dt = _DT_NUMPY[e.dtype]
        n = int(np.prod(e.shape)) if e.shape else 1
        if n * dt.itemsize != e.size:
            raise CheckpointError(f"{key}: size/shape mismatch")
        out[key] = np.frombuffer(raw, dtype=dt).reshape(e.shape).copy()
    return out


_SUFFIX = "/.ATTRIBUTES/VARIABLE_VALUE"


def load_net_weights(model_dir_or_prefix: str) -> List[Dict[str, np.ndarray]]:
    """Return the `net/layer_with_weights-k` tensors as an ordered list.

    Even k are Conv2D {kernel[HWIO], bias}; odd k are BatchNormalization
    {gamma, beta, moving_mean, moving_variance} (Keras functional order of
    /root/reference/models/model.py:30-41 and :5-28).
    """
    prefix = model_dir_or_prefix
    if os.path.isdir(prefix):
        latest = latest_checkpoint(prefix)
        if latest is None:
            raise CheckpointError(f"no checkpoint state file in {prefix}")
        prefix = latest
    tensors = read_bundle(prefix, keys_filter=lambda k: k.startswith("net/") and
                          k.endswith(_SUFFIX) and ".OPTIMIZER_SLOT" not in k)
    layers: Dict[int, Dict[str, np.ndarray]] = {}
    for key, arr in tensors.items():
        parts = key[:-len(_SUFFIX)].split("/")
        if len(parts) != 3 or not parts[1].startswith("layer_with_weights-"):
            continue
        k = int(parts[1].split("-")[1])
        layers.setdefault(k, {})[parts[2]] = arr
    if not layers:
        raise CheckpointError(f"{prefix}: no net/layer_with_weights-* tensors")
    return [layers[k] for k in sorted(layers)]


# ---- writing ---------------------------------------------------------------
def _block(entries: List[Tuple[bytes, bytes]]) -> bytes:
    # one restart point per entry: no prefix compression, trivially valid
    body = bytearray()
    restarts = []
    for key, val in entries:
        restarts.append(len(body))
        body += _put_varint(0) + _put_varint(len(key)) + _put_varint(len(val)) + key + val
    for r in restarts:
        body += struct.pack("<I", r)
    body += struct.pack("<I", len(restarts))
    return bytes(body)


def _with_trailer(block: bytes) -> bytes:
    crc = masked_crc32c(block + b"\x00")
    return block + b"\x00" + struct.pack("<I", crc)


def _entry_proto(dtype: int, shape, offset: int, size: int, crc: int) -> bytes:
    dims = b""
    for d in shape:
        dim = b"\x08" + _put_varint(int(d))
        dims += b"\x12" + _put_varint(len(dim)) + dim
    out = b"\x08" + _put_varint(dtype)
    out += b"\x12" + _put_varint(len(dims)) + dims
    if offset:
        out += b"\x20" + _put_varint(offset)
    out += b"\x28" + _put_varint(size)
    out += b"\x35" + struct.pack("<I", crc)
    return out


def write_bundle(prefix: str, tensors: Dict[str, np.ndarray]) -> None:
    """Write {key: ndarray | bytes} as a single-shard TensorBundle (`bytes` -> scalar string tensor)."""
    data = bytearray()
    entries: List[Tuple[bytes, bytes]] = []
    # BundleHeaderProto{num_shards=1, endianness=LITTLE(0, omitted), version{producer=1}}
    entries.append((b"", b"\x08\x01\x1a\x02\x08\x01"))
    for key in sorted(tensors):
        if isinstance(tensors[key], (bytes, bytearray)):
            raw, crc = _string_payload(bytes(tensors[key]))
            entries.append((key.encode("utf-8"), _entry_proto(_DT_STRING, (), len(data), len(raw), crc)))
            data += raw
            continue
        arr = np.asarray(tensors[key], order="C")
        dt = arr.dtype.newbyteorder("<") if arr.dtype.byteorder == ">" else arr.dtype
        if np.dtype(dt) not in _NUMPY_DT:
            raise CheckpointError(f"{key}: unsupported dtype {arr.dtype}")
        raw = arr.astype(dt, copy=False).tobytes()
        entries.append((key.encode("utf-8"),
                        _entry_proto(_NUMPY_DT[np.dtype(dt)], arr.shape, len(data),
                                     len(raw), masked_crc32c(raw))))
        data += raw
    data_block = _block(entries)
    out = bytearray(_with_trailer(data_block))
    meta_off = len(out)
    meta_block = _block([])
    out += _with_trailer(meta_block)
    idx_off = len(out)
    handle = _put_varint(0) + _put_varint(len(data_block))
    idx_block = _block([(entries[-1][0] + b"\xff", handle)])
    out += _with_trailer(idx_block)
    footer = (_put_varint(meta_off) + _put_varint(len(meta_block)) +
              _put_varint(idx_off) + _put_varint(len(idx_block)))
    footer += b"\x00" * (40 - len(footer)) + struct.pack("<Q", _MAGIC)
    out += footer
    with open(prefix + ".index", "wb") as f:
        f.write(bytes(out))
    with open(prefix + ".data-00000-of-00001", "wb") as f:
        f.write(bytes(data))


# ---- the object graph of tf.train.Checkpoint(step, optimizer, net) ---------------------------------------------
OBJECT_GRAPH_KEY = "_CHECKPOINTABLE_OBJECT_GRAPH"
_CONV_VARS = ("kernel", "bias")
_BN_VARS = ("gamma", "beta", "moving_mean", "moving_variance")
_ADAM_HYPERS = ("iter", "beta_1", "beta_2", "decay", "learning_rate")


def _pb_str(field: int, value: bytes) -> bytes:
    return _put_varint(field << 3 | 2) + _put_varint(len(value)) + value


def _pb_int(field: int, value: int) -> bytes:
    return _put_varint(field << 3) + _put_varint(value) if value else b""      # proto3: zero is not written


def _keras_name(base: str, index: int) -> str:
    return base if index == 0 else "%s_%d" % (base, index)


def object_graph(n_blocks: int, conv_names=None, bn_names=None) -> bytes:
    """Serialised `TrackableObjectGraph` (tensorflow/core/protobuf/trackable_object_graph.proto) of the
    reference's `tf.train.Checkpoint(step=..., optimizer=Adam, net=model)` (train.py:78-80) where `model` is a
    functional Keras model of `n_blocks` Conv2D + BatchNormalization pairs (models/model.py).

    Node numbering and field order follow what TensorFlow wrote into the shipped ckpt-1000: root; net, optimizer,
    step, save_counter; the InputLayer and the 2*n_blocks layers; the optimizer's hyper-parameters; the layers'
    variables (a BatchNormalization also tracks an `axis` list wrapper, a node without attributes); then one Adam
    `m` slot per trainable variable, then the `v` slots.  `conv_names` / `bn_names` are the Keras layer names
    (`full_name` is informational; restore matches on the graph)."""
    conv_names = conv_names or [_keras_name("conv2d", i) for i in range(n_blocks)]
    bn_names = bn_names or [_keras_name("batch_normalization", i) for i in range(n_blocks)]
    n_layers = 2 * n_blocks
    nodes = []          # each: {"children": [(id, name)], "attr": (full_name, key) | None, "slots": [(orig, slot, id)]}

    def new_node():
        nodes.append({"children": [], "attr": None, "slots": []})
        return len(nodes) - 1

    root, net, opt, step, counter = (new_node() for _ in range(5))
    nodes[root]["children"] = [(net, "net"), (opt, "optimizer"), (step, "step"), (counter, "save_counter")]
    nodes[step]["attr"] = ("Variable", "step" + _SUFFIX)
    nodes[counter]["attr"] = ("save_counter", "save_counter" + _SUFFIX)
    input_layer = new_node()
    layer_nodes = [new_node() for _ in range(n_layers)]
    nodes[net]["children"].append((input_layer, "layer-0"))
    for k, node in enumerate(layer_nodes):
        nodes[net]["children"] += [(node, "layer_with_weights-%d" % k), (node, "layer-%d" % (k + 1))]
    for name in _ADAM_HYPERS:
        node = new_node()
        nodes[opt]["children"].append((node, name))
        nodes[node]["attr"] = ("Adam/" + name, "optimizer/%s%s" % (name, _SUFFIX))
    trainable = []      # (node id, full name, checkpoint key stem)
    for k, layer in enumerate(layer_nodes):
        is_conv = k % 2 == 0
        lname = conv_names[k // 2] if is_conv else bn_names[k // 2]
        if not is_conv:
            nodes[layer]["children"].append((new_node(), "axis"))
        for var in (_CONV_VARS if is_conv else _BN_VARS):
            node = new_node()
            nodes[layer]["children"].append((node, var))
            stem = "net/layer_with_weights-%d/%s" % (k, var)
            nodes[node]["attr"] = ("%s/%s" % (lname, var), stem + _SUFFIX)
            if not var.startswith("moving_"):
                trainable.append((node, "%s/%s" % (lname, var), stem))
    for slot in ("m", "v"):
        for orig, full, stem in trainable:
            node = new_node()
            nodes[node]["attr"] = ("Adam/%s/%s" % (full, slot), "%s/.OPTIMIZER_SLOT/optimizer/%s%s" % (stem, slot, _SUFFIX))
            nodes[opt]["slots"].append((orig, slot, node))

    out = b""
    for nd in nodes:
        body = b""
        for cid, name in nd["children"]:
            body += _pb_str(1, _pb_int(1, cid) + _pb_str(2, name.encode()))
        if nd["attr"]:
            full, key = nd["attr"]
            body += _pb_str(2, _pb_str(1, b"VARIABLE_VALUE") + _pb_str(2, full.encode()) + _pb_str(3, key.encode()))
        for orig, slot, sid in nd["slots"]:
            body += _pb_str(3, _pb_int(1, orig) + _pb_str(2, slot.encode()) + _pb_int(3, sid))
        out += _pb_str(1, body)
    return out
