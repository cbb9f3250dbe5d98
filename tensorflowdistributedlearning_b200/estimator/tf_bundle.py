"""TensorFlow "tensor bundle" checkpoints without TensorFlow: ``<prefix>.index`` + ``<prefix>.data-00000-of-00001``.

The reference saves its variables with ``tf.train.Saver`` (V2 format) through the Estimator
(model.py:117-121,166).  This module reads such checkpoints and writes files of the same
layout, so weights trained with the reference can be loaded here and vice versa:

* ``.index`` is a LevelDB-style sorted string table: data blocks of prefix-compressed (key, value) entries with a
  restart array, a 5-byte trailer per block (compression type + masked CRC-32C), a metaindex block, an index block
  mapping separator keys to block handles, and a 48-byte footer ending in the magic ``0xdb4775248b80fb57``.
  The empty key holds a ``BundleHeaderProto`` (num_shards, endianness, version); every other key is a variable name
  whose value is a ``BundleEntryProto`` (dtype, shape, shard_id, offset, size, crc32c).
* ``.data-*`` holds the raw little-endian tensor bytes at those offsets.

Only what those two messages need of protobuf is implemented (varints, length-delimited fields, fixed32).
CRC-32C comes from the native library (``tdl_crc32c``), with a slow pure-Python fallback.
"""
from __future__ import annotations

import ctypes
import os
import struct
from typing import Dict, Iterable, List, Tuple

import numpy as np

MAGIC = 0xDB4775248B80FB57
_MASK_DELTA = 0xA282EAD8

# tensorflow/core/framework/types.proto
_DT = {1: np.float32, 2: np.float64, 3: np.int32, 4: np.uint8, 5: np.int16, 6: np.int8, 9: np.int64, 10: np.bool_,
       17: np.uint16, 19: np.float16, 22: np.uint32, 23: np.uint64}
_DT_BFLOAT16 = 14
_NP2DT = {np.dtype(v): k for k, v in _DT.items()}


# ---- crc32c -----------------------------------------------------------------------------------------------
_py_table = None


def _crc32c_py(data: bytes, crc: int = 0) -> int:
    global _py_table
    if _py_table is None:
        tab = []
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
            tab.append(c)
        _py_table = tab
    c = crc ^ 0xFFFFFFFF
    for b in data:
        c = (c >> 8) ^ _py_table[(c ^ b) & 0xFF]
    return c ^ 0xFFFFFFFF


def crc32c(data, crc: int = 0) -> int:
    """CRC-32C (Castagnoli) of a bytes-like object."""
    buf = bytes(data) if not isinstance(data, (bytes, bytearray)) else data
    try:
        from .. import _native as N

        fn = N.lib().tdl_crc32c
        fn.restype = ctypes.c_uint32
        return int(fn(ctypes.c_char_p(bytes(buf)), ctypes.c_uint64(len(buf)), ctypes.c_uint32(crc)))
    except Exception:  # noqa: BLE001 -- library not built: correct but slow
        return _crc32c_py(bytes(buf), crc)


def masked_crc(data) -> int:
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + _MASK_DELTA) & 0xFFFFFFFF


def unmask_crc(m: int) -> int:
    rot = (m - _MASK_DELTA) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


# ---- varints / minimal protobuf -------------------------------------------------------------------------
def _put_varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _get_varint(buf: bytes, pos: int) -> Tuple[int, int]:
    shift = n = 0
    while True:
        b = buf[pos]
        pos += 1
        n |= (b & 0x7F) << shift
        if not b & 0x80:
            return n, pos
        shift += 7


def _pb_fields(buf: bytes) -> Iterable[Tuple[int, int, object]]:
    """(field number, wire type, value) for every field of a serialized message."""
    pos = 0
    while pos < len(buf):
        key, pos = _get_varint(buf, pos)
        field, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _get_varint(buf, pos)
        elif wt == 1:
            v = buf[pos:pos + 8]
            pos += 8
        elif wt == 2:
            ln, pos = _get_varint(buf, pos)
            v = buf[pos:pos + ln]
            pos += ln
        elif wt == 5:
            v = struct.unpack_from("<I", buf, pos)[0]
            pos += 4
        else:
            raise ValueError(f"unsupported protobuf wire type {wt}")
        yield field, wt, v


def _pb_varint_field(field: int, v: int) -> bytes:
    return _put_varint(field << 3) + _put_varint(v)


def _pb_bytes_field(field: int, b: bytes) -> bytes:
    return _put_varint((field << 3) | 2) + _put_varint(len(b)) + b


def _encode_shape(shape: Tuple[int, ...]) -> bytes:
    return b"".join(_pb_bytes_field(2, _pb_varint_field(1, int(d))) for d in shape)  # TensorShapeProto.dim[].size


def _decode_shape(buf: bytes) -> Tuple[int, ...]:
    dims = []
    for f, wt, v in _pb_fields(buf):
        if f == 2 and wt == 2:
            size = 0
            for f2, _, v2 in _pb_fields(v):
                if f2 == 1:
                    size = v2
            dims.append(int(size))
    return tuple(dims)


def _encode_entry(dtype: int, shape, offset: int, size: int, crc: int) -> bytes:
    out = _pb_varint_field(1, dtype) + _pb_bytes_field(2, _encode_shape(shape))
    if offset:
        out += _pb_varint_field(4, offset)
    out += _pb_varint_field(5, size)
    out += _put_varint((6 << 3) | 5) + struct.pack("<I", crc)
    return out


def _decode_entry(buf: bytes) -> dict:
    e = {"dtype": 0, "shape": (), "shard_id": 0, "offset": 0, "size": 0, "crc32c": None, "slices": False}
    for f, wt, v in _pb_fields(buf):
        if f == 1:
            e["dtype"] = v
        elif f == 2:
            e["shape"] = _decode_shape(v)
        elif f == 3:
            e["shard_id"] = v
        elif f == 4:
            e["offset"] = v
        elif f == 5:
            e["size"] = v
        elif f == 6:
            e["crc32c"] = v
        elif f == 7:
            e["slices"] = True
    return e


# ---- sorted string table ------------------------------------------------------------------------------------
def _block_entries(block: bytes) -> List[Tuple[bytes, bytes]]:
    num_restarts = struct.unpack_from("<I", block, len(block) - 4)[0]
    end = len(block) - 4 - 4 * num_restarts
    pos, key, out = 0, b"", []
    while pos < end:
        shared, pos = _get_varint(block, pos)
        unshared, pos = _get_varint(block, pos)
        vlen, pos = _get_varint(block, pos)
        key = key[:shared] + block[pos:pos + unshared]
        pos += unshared
        out.append((key, block[pos:pos + vlen]))
        pos += vlen
    return out


def _read_block(f, offset: int, size: int, verify: bool = True) -> bytes:
    f.seek(offset)
    raw = f.read(size + 5)
    block, ctype, crc = raw[:size], raw[size], struct.unpack_from("<I", raw, size + 1)[0]
    if verify and unmask_crc(crc) != crc32c(raw[:size + 1]):
        raise IOError("tensor bundle index: block checksum mismatch")
    if ctype != 0:
        raise NotImplementedError("compressed index blocks (snappy) are not supported; TF writes bundles uncompressed")
    return block


def read_index(prefix: str) -> Tuple[dict, Dict[str, dict]]:
    """-> (header, {variable name: entry dict})"""
    path = prefix + ".index"
    size = os.path.getsize(path)
    with open(path, "rb") as f:
        f.seek(size - 48)
        footer = f.read(48)
        if struct.unpack_from("<Q", footer, 40)[0] != MAGIC:
            raise IOError(f"{path} is not a TensorFlow table (bad magic)")
        pos = 0
        _, pos = _get_varint(footer, pos)  # metaindex handle (unused)
        _, pos = _get_varint(footer, pos)
        idx_off, pos = _get_varint(footer, pos)
        idx_size, pos = _get_varint(footer, pos)
        header, entries = {}, {}
        for _, handle in _block_entries(_read_block(f, idx_off, idx_size)):
            off, p2 = _get_varint(handle, 0)
            sz, _ = _get_varint(handle, p2)
            for key, value in _block_entries(_read_block(f, off, sz)):
                if key == b"":
                    for fld, _, v in _pb_fields(value):
                        if fld == 1:
                            header["num_shards"] = v
                        elif fld == 2:
                            header["endianness"] = v
                    header.setdefault("num_shards", 1)
                    header.setdefault("endianness", 0)
                else:
                    entries[key.decode()] = _decode_entry(value)
    return header, entries


def read_bundle(prefix: str, verify: bool = True) -> Dict[str, np.ndarray]:
    """Every tensor of a TF V2 checkpoint as a NumPy array (bfloat16 comes back as float32)."""
    header, entries = read_index(prefix)
    if header.get("endianness", 0) != 0:
        raise NotImplementedError("big-endian bundles are not supported")
    n = header.get("num_shards", 1)
    files, out = {}, {}
    try:
        for name, e in entries.items():
            if e["slices"]:
                raise NotImplementedError(f"{name}: partitioned variables (tensor slices) are not supported")
            sid = e["shard_id"]
            if sid not in files:
                files[sid] = open(f"{prefix}.data-{sid:05d}-of-{n:05d}", "rb")
            f = files[sid]
            f.seek(e["offset"])
            raw = f.read(e["size"])
            if verify and e["crc32c"] is not None and unmask_crc(e["crc32c"]) != crc32c(raw):
                raise IOError(f"{name}: tensor checksum mismatch")
            if e["dtype"] == _DT_BFLOAT16:
                arr = (np.frombuffer(raw, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)
            elif e["dtype"] in _DT:
                arr = np.frombuffer(raw, dtype=_DT[e["dtype"]])
            else:
                raise NotImplementedError(f"{name}: unsupported dtype enum {e['dtype']}")
            out[name] = arr.reshape(e["shape"]).copy()
    finally:
        for f in files.values():
            f.close()
    return out


class _BlockBuilder:
    """One table block; every entry is its own restart point (shared prefix 0), which every reader accepts."""

    def __init__(self):
        self.buf, self.restarts, self.last_key = bytearray(), [], b""

    def add(self, key: bytes, value: bytes) -> None:
        self.restarts.append(len(self.buf))
        self.buf += _put_varint(0) + _put_varint(len(key)) + _put_varint(len(value)) + key + value
        self.last_key = key

    def finish(self) -> bytes:
        restarts = self.restarts or [0]
        return bytes(self.buf) + b"".join(struct.pack("<I", r) for r in restarts) + struct.pack("<I", len(restarts))


def _write_block(f, content: bytes) -> bytes:
    """Writes block + trailer, returns its encoded BlockHandle."""
    off = f.tell()
    trailer_type = b"\x00"
    f.write(content + trailer_type + struct.pack("<I", masked_crc(content + trailer_type)))
    return _put_varint(off) + _put_varint(len(content))


def write_bundle(prefix: str, tensors: Dict[str, np.ndarray], block_size: int = 4096) -> None:
    """Write ``{name: array}`` as ``<prefix>.index`` + ``<prefix>.data-00000-of-00001`` (one shard, little endian)."""
    names = sorted(tensors, key=lambda s: s.encode())
    entries = []
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    with open(prefix + ".data-00000-of-00001.tmp", "wb") as f:
        for name in names:
            a = np.asarray(tensors[name])
            a = np.ascontiguousarray(a).reshape(a.shape)  # (ascontiguousarray alone turns scalars into shape (1,))
            if a.dtype not in _NP2DT:
                raise TypeError(f"{name}: dtype {a.dtype} has no TensorFlow equivalent here")
            if a.dtype.byteorder == ">":
                a = a.astype(a.dtype.newbyteorder("<"))
            raw = a.tobytes()
            entries.append((name.encode(), _encode_entry(_NP2DT[a.dtype], a.shape, f.tell(), len(raw), masked_crc(raw))))
            f.write(raw)
    os.replace(prefix + ".data-00000-of-00001.tmp", prefix + ".data-00000-of-00001")
    header = _pb_varint_field(1, 1) + _pb_bytes_field(3, _pb_varint_field(1, 1))  # num_shards=1, version.producer=1
    with open(prefix + ".index.tmp", "wb") as f:
        index = _BlockBuilder()
        block = _BlockBuilder()
        for key, value in [(b"", header)] + entries:
            block.add(key, value)
            if len(block.buf) >= block_size:
                index.add(block.last_key, _write_block(f, block.finish()))
                block = _BlockBuilder()
        if block.restarts:
            index.add(block.last_key, _write_block(f, block.finish()))
        meta_handle = _write_block(f, _BlockBuilder().finish())
        index_handle = _write_block(f, index.finish())
        footer = meta_handle + index_handle
        f.write(footer + b"\x00" * (40 - len(footer)) + struct.pack("<Q", MAGIC))
    os.replace(prefix + ".index.tmp", prefix + ".index")
