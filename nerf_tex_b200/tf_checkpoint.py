"""Weights in: reader for the TensorFlow checkpoints the reference's Logger writes
(network/logger.py:30-40: ``tf.train.Checkpoint(model=..., model_fine=..., step=..., optimizer=...)`` managed by a
``CheckpointManager`` in ``<target>/checkpoints``), without TensorFlow.

Format (TensorFlow's "tensor bundle"):
  * ``checkpoint``                      text file, first line ``model_checkpoint_path: "ckpt-<n>"``
  * ``ckpt-<n>.index``                  a LevelDB-format sorted string table: key "" -> BundleHeaderProto, every other
                                        key (``model/layer_with_weights-3/kernel/.ATTRIBUTES/VARIABLE_VALUE`` ...) ->
                                        BundleEntryProto {dtype, shape, shard_id, offset, size, crc32c}
  * ``ckpt-<n>.data-00000-of-00001``    raw little-endian tensor bytes at those offsets
The table reader handles prefix-compressed keys, restart arrays, multi-block tables and both block encodings
(raw and snappy); block and tensor checksums are masked CRC-32C.

STATUS: written from the published format description.  Neither TensorFlow nor a TensorFlow-written checkpoint
exists in this environment (the reference repo ships none), so the reader is only tested against files produced by
the ``write_bundle`` function below, which follows the same description — a self-consistency check, not proof of
interoperability."""

from __future__ import annotations

import os
import re
import struct
from typing import Dict, Iterable, List, NamedTuple, Optional, Tuple

import numpy as np

TABLE_MAGIC = 0xdb4775248b80fb57
DTYPES = {1: np.float32, 2: np.float64, 3: np.int32, 4: np.uint8, 5: np.int16, 6: np.int8, 9: np.int64, 10: np.bool_,
          17: np.uint16, 19: np.float16, 22: np.uint32, 23: np.uint64}
DT_OF = {np.dtype(v): k for k, v in DTYPES.items()}
VALUE_SUFFIX = "/.ATTRIBUTES/VARIABLE_VALUE"


# ---------------------------------------------------------------------------------------------- checksums, varints
def _crc_table() -> List[int]:
    tab = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
        tab.append(c)
    return tab


_CRC = _crc_table()


def crc32c(data: bytes, crc: int = 0) -> int:
    crc ^= 0xFFFFFFFF
    for b in data:
        crc = _CRC[(crc ^ b) & 0xFF] ^ (crc >> 8)
    return crc ^ 0xFFFFFFFF


def masked_crc(data: bytes) -> int:
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xa282ead8) & 0xFFFFFFFF


def _varint(buf: bytes, pos: int) -> Tuple[int, int]:
    out = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if b < 0x80:
            return out, pos
        shift += 7


def _put_varint(v: int) -> bytes:
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


# ---------------------------------------------------------------------------------------------- protobuf (tiny subset)
def _proto_fields(buf: bytes) -> Iterable[Tuple[int, int, object]]:
    pos = 0
    while pos < len(buf):
        tag, pos = _varint(buf, pos)
        field, wire = tag >> 3, tag & 7
        if wire == 0:
            v, pos = _varint(buf, pos)
        elif wire == 1:
            v, pos = buf[pos:pos + 8], pos + 8
        elif wire == 2:
            n, pos = _varint(buf, pos)
            v, pos = buf[pos:pos + n], pos + n
        elif wire == 5:
            v, pos = buf[pos:pos + 4], pos + 4
        else:
            raise ValueError(f"unsupported protobuf wire type {wire}")
        yield field, wire, v


class Entry(NamedTuple):
    dtype: int
    shape: Tuple[int, ...]
    shard_id: int
    offset: int
    size: int
    crc32c: Optional[int]


def _signed64(v: int) -> int:
    return v - (1 << 64) if v >= (1 << 63) else v


def _parse_entry(buf: bytes) -> Entry:
    dtype = shard = offset = size = 0
    shape: List[int] = []
    crc = None
    for f, w, v in _proto_fields(buf):
        if f == 1:
            dtype = v
        elif f == 2:                                    # TensorShapeProto { repeated Dim dim = 2 { int64 size = 1 } }
            for f2, _, v2 in _proto_fields(v):
                if f2 == 2:
                    dim = 0
                    for f3, _, v3 in _proto_fields(v2):
                        if f3 == 1:
                            dim = _signed64(v3)
                    shape.append(dim)
        elif f == 3:
            shard = v
        elif f == 4:
            offset = v
        elif f == 5:
            size = v
        elif f == 6:
            crc = struct.unpack("<I", v)[0]
        elif f == 7:
            raise NotImplementedError("sliced (partitioned) variables are not supported")
    return Entry(dtype, tuple(shape), shard, offset, size, crc)


def _encode_entry(e: Entry) -> bytes:
    dims = b"".join(b"\x12" + _put_varint(len(d)) + d for d in (b"\x08" + _put_varint(s) for s in e.shape))
    out = b"\x08" + _put_varint(e.dtype) + b"\x12" + _put_varint(len(dims)) + dims
    if e.shard_id:
        out += b"\x18" + _put_varint(e.shard_id)
    if e.offset:
        out += b"\x20" + _put_varint(e.offset)
    out += b"\x28" + _put_varint(e.size)
    if e.crc32c is not None:
        out += b"\x35" + struct.pack("<I", e.crc32c)
    return out


# ---------------------------------------------------------------------------------------------- snappy (decode only)
def _snappy_decompress(buf: bytes) -> bytes:
    n, pos = _varint(buf, 0)
    out = bytearray()
    while pos < len(buf):
        tag = buf[pos]
        pos += 1
        kind = tag & 3
        if kind == 0:                                   # literal
            ln = tag >> 2
            if ln >= 60:
                nb = ln - 59
                ln = int.from_bytes(buf[pos:pos + nb], "little")
                pos += nb
            ln += 1
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
        if off == 0 or off > len(out):
            raise ValueError("corrupt snappy stream")
        for _ in range(ln):                             # may overlap its own output
            out.append(out[-off])
    if len(out) != n:
        raise ValueError("corrupt snappy stream (length)")
    return bytes(out)


# ---------------------------------------------------------------------------------------------- sorted string table
def _read_block(data: bytes, offset: int, size: int, verify: bool = True) -> bytes:
    body, ctype = data[offset:offset + size], data[offset + size]
    if verify:
        want = struct.unpack("<I", data[offset + size + 1:offset + size + 5])[0]
        if masked_crc(data[offset:offset + size + 1]) != want:
            raise ValueError(f"index block at {offset}: checksum mismatch")
    if ctype == 0:
        return body
    if ctype == 1:
        return _snappy_decompress(body)
    raise ValueError(f"index block at {offset}: unknown compression type {ctype}")


def _block_entries(block: bytes) -> Iterable[Tuple[bytes, bytes]]:
    n_restarts = struct.unpack("<I", block[-4:])[0]
    end = len(block) - 4 - 4 * n_restarts
    pos, key = 0, b""
    while pos < end:
        shared, pos = _varint(block, pos)
        non_shared, pos = _varint(block, pos)
        vlen, pos = _varint(block, pos)
        key = key[:shared] + block[pos:pos + non_shared]
        pos += non_shared
        yield key, block[pos:pos + vlen]
        pos += vlen


def read_table(path: str, verify: bool = True) -> Dict[bytes, bytes]:
    data = open(path, "rb").read()
    if len(data) < 48 or struct.unpack("<Q", data[-8:])[0] != TABLE_MAGIC:
        raise ValueError(f"{path}: not a sorted string table (bad magic)")
    footer = data[-48:]
    _, p = _varint(footer, 0)                           # metaindex handle (unused)
    _, p = _varint(footer, p)
    ioff, p = _varint(footer, p)
    isize, p = _varint(footer, p)
    out: Dict[bytes, bytes] = {}
    for _, handle in _block_entries(_read_block(data, ioff, isize, verify)):
        boff, q = _varint(handle, 0)
        bsize, _ = _varint(handle, q)
        for k, v in _block_entries(_read_block(data, boff, bsize, verify)):
            out[k] = v
    return out


# ---------------------------------------------------------------------------------------------- bundle
def read_index(prefix: str, verify: bool = True) -> Tuple[int, Dict[str, Entry]]:
    """(num_shards, {tensor key: Entry})."""
    table = read_table(prefix + ".index", verify)
    if b"" not in table:
        raise ValueError(f"{prefix}.index: no bundle header")
    num_shards, endian = 1, 0
    for f, _, v in _proto_fields(table[b""]):
        if f == 1:
            num_shards = v
        elif f == 2:
            endian = v
    if endian != 0:
        raise NotImplementedError("big-endian checkpoints are not supported")
    return num_shards, {k.decode(): _parse_entry(v) for k, v in table.items() if k != b""}


def read_tensor(prefix: str, entry: Entry, num_shards: int = 1, verify: bool = False) -> np.ndarray:
    if entry.dtype not in DTYPES:
        raise NotImplementedError(f"tensor dtype enum {entry.dtype} is not supported")
    path = "{}.data-{:05d}-of-{:05d}".format(prefix, entry.shard_id, num_shards)
    with open(path, "rb") as f:
        f.seek(entry.offset)
        raw = f.read(entry.size)
    if len(raw) != entry.size:
        raise ValueError(f"{path}: truncated")
    if verify and entry.crc32c is not None and masked_crc(raw) != entry.crc32c:
        raise ValueError(f"{path} @ {entry.offset}: tensor checksum mismatch")
    arr = np.frombuffer(raw, dtype=DTYPES[entry.dtype])
    return arr.reshape(entry.shape).copy()


def latest_checkpoint(directory: str) -> Optional[str]:
    """Prefix named by the ``checkpoint`` state file (tf.train.latest_checkpoint)."""
    state = os.path.join(directory, "checkpoint")
    if not os.path.exists(state):
        return None
    for line in open(state):
        m = re.match(r'\s*model_checkpoint_path:\s*"(.*)"', line)
        if m:
            p = m.group(1)
            return p if os.path.isabs(p) else os.path.join(directory, p)
    return None


def dense_layers(index: Dict[str, Entry], name: str) -> List[Tuple[str, str]]:
    """[(kernel key, bias key)] of the Keras model saved under attribute `name`, in layer_with_weights order."""
    pat = re.compile(re.escape(name) + r"/layer_with_weights-(\d+)/kernel" + re.escape(VALUE_SUFFIX) + "$")
    found = sorted((int(m.group(1)), k) for k in index for m in [pat.match(k)] if m)
    return [(k, k.replace("/kernel/", "/bias/")) for _, k in found]


def load_model_weights(prefix: str, model, name: str, verify: bool = False) -> Dict[str, Tuple[np.ndarray, np.ndarray]]:
    """Fill a FusedNerfTexMLP from the Keras model stored as `name` ('model' / 'model_fine', network/model.py:45,125).

    Keras numbers the weighted layers of a functional model (layer_with_weights-i) in its own traversal order, in which
    the density head may come before or after the colour branch.  Layers are therefore matched by kernel shape, and
    layers of equal shape by their order: trunk layers are created in order, and the feature layer after the last
    trunk layer in either numbering (network/model.py:104-123)."""
    num_shards, index = read_index(prefix)
    pairs = dense_layers(index, name)
    want = list(zip(model.layer_names, [(fi, fo) for _, _, fi, fo in model.layers]))
    if len(pairs) != len(want):
        raise ValueError(f"{prefix}: '{name}' has {len(pairs)} Dense layers, the model {len(want)}")
    by_shape: Dict[Tuple[int, int], List[Tuple[str, str]]] = {}
    for kk, bk in pairs:
        by_shape.setdefault(tuple(index[kk].shape), []).append((kk, bk))
    named = {}
    for lname, shape in want:                            # model order within a shape class == creation order
        bucket = by_shape.get(shape)
        if not bucket:
            raise ValueError(f"{prefix}: no kernel of shape {shape} left for layer {lname}")
        kk, bk = bucket.pop(0)
        named[lname] = (read_tensor(prefix, index[kk], num_shards, verify), read_tensor(prefix, index[bk], num_shards, verify))
    model.load_named_weights(named)
    return named


def read_step(prefix: str) -> Optional[int]:
    num_shards, index = read_index(prefix)
    e = index.get("step" + VALUE_SUFFIX)
    return None if e is None else int(read_tensor(prefix, e, num_shards))


# ---------------------------------------------------------------------------------------------- writer (tests, export)
def write_bundle(prefix: str, tensors: Dict[str, np.ndarray], block_size: int = 4096, restart_interval: int = 16) -> None:
    """Write {key: array} as a single-shard bundle (uncompressed blocks), plus nothing else (no ``checkpoint`` file)."""
    os.makedirs(os.path.dirname(prefix) or ".", exist_ok=True)
    entries: Dict[bytes, bytes] = {b"": b"\x08\x01\x1a\x02\x08\x01"}        # num_shards 1, little endian, version {producer 1}
    offset = 0
    with open(prefix + ".data-00000-of-00001", "wb") as f:
        for key in sorted(tensors):
            a = np.asarray(tensors[key])
            a = a if a.flags.c_contiguous else np.ascontiguousarray(a)     # (ascontiguousarray turns 0-d into 1-d)
            raw = a.tobytes()
            f.write(raw)
            entries[key.encode()] = _encode_entry(Entry(DT_OF[a.dtype], tuple(a.shape), 0, offset, len(raw), masked_crc(raw)))
            offset += len(raw)

    out = bytearray()

    def emit_block(items: List[Tuple[bytes, bytes]]) -> bytes:
        body, restarts, prev = bytearray(), [], b""
        for i, (k, v) in enumerate(items):
            shared = 0
            if i % restart_interval == 0:
                restarts.append(len(body))
            else:
                while shared < min(len(prev), len(k)) and prev[shared] == k[shared]:
                    shared += 1
            body += _put_varint(shared) + _put_varint(len(k) - shared) + _put_varint(len(v)) + k[shared:] + v
            prev = k
        body += b"".join(struct.pack("<I", r) for r in (restarts or [0])) + struct.pack("<I", max(len(restarts), 1))
        handle = _put_varint(len(out)) + _put_varint(len(body))
        out.extend(body + b"\x00" + struct.pack("<I", masked_crc(bytes(body) + b"\x00")))
        return handle

    index_items, block, size = [], [], 0
    for k in sorted(entries):
        block.append((k, entries[k]))
        size += len(k) + len(entries[k])
        if size >= block_size:
            index_items.append((block[-1][0], emit_block(block)))
            block, size = [], 0
    if block:
        index_items.append((block[-1][0], emit_block(block)))
    meta = emit_block([])
    idx = emit_block(index_items)
    footer = meta + idx
    out.extend(footer + b"\x00" * (40 - len(footer)) + struct.pack("<Q", TABLE_MAGIC))
    with open(prefix + ".index", "wb") as f:
        f.write(bytes(out))
