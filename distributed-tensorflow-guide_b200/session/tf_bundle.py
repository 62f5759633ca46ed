"""TensorFlow checkpoint V2 ("tensor bundle") reader / writer, in pure Python.

SURVEY 2.6 keeps the reference's checkpoint *contract* (directory layout, the ``checkpoint``
state file, TF variable names as keys) with our own ``.dtg`` payload; this module is the
stretch item next to it: the on-disk format TF's Saver V2 itself produces, so a ``logdir``
written by the reference scripts (``model.ckpt-<step>.index`` + ``.data-00000-of-00001``)
can be read here and a checkpoint written here has the layout TF expects.

Format (tensorflow/core/util/tensor_bundle, tensorflow/core/lib/io/table_*; no TF is available
offline, so interoperability is by construction from that specification and verified here by
round trips and structural checks only):

* ``<prefix>.data-00000-of-00001``: the tensors' raw little-endian bytes, concatenated.
* ``<prefix>.index``: a leveldb-style sorted string table.  Data blocks hold prefix-compressed
  ``key -> value`` entries (restart interval 16) and end with the restart array; every block
  is followed by a 5-byte trailer (compression type 0, masked CRC32C).  An (empty) meta-index
  block, an index block (last key of each data block -> block handle) and a 48-byte footer
  (two block handles, padding, magic ``0xdb4775248b80fb57``) close the file.
* key ``""`` -> ``BundleHeaderProto{num_shards=1, endianness=LITTLE, version{producer=1}}``;
  key ``<variable name>`` -> ``BundleEntryProto{dtype, shape, shard_id, offset, size, crc32c}``
  where ``crc32c`` is the masked CRC32C of the tensor's bytes.
"""
from __future__ import annotations

import os
import struct
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np
import torch

__all__ = ["write_bundle", "read_bundle", "list_bundle", "is_bundle", "crc32c", "masked_crc32c"]

_MAGIC = 0xdb4775248b80fb57
_RESTART = 16
_BLOCK_SIZE = 4096

# tensorflow/core/framework/types.proto
_DT = {torch.float32: 1, torch.float64: 2, torch.int32: 3, torch.uint8: 4, torch.int16: 5,
       torch.int8: 6, torch.int64: 9, torch.bool: 10, torch.bfloat16: 14, torch.float16: 19}
_DT_INV = {v: k for k, v in _DT.items()}
_NP = {torch.float32: "<f4", torch.float64: "<f8", torch.int32: "<i4", torch.uint8: "|u1",
       torch.int16: "<i2", torch.int8: "|i1", torch.int64: "<i8", torch.bool: "|b1",
       torch.float16: "<f2"}


# ------------------------------------------------------------------------- crc32c
def _make_table() -> List[int]:
    poly, table = 0x82F63B78, []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ poly if c & 1 else c >> 1
        table.append(c)
    return table


_TABLE = _make_table()


_native = None


def _native_crc():
    """``dtg_crc32c`` of the native library (slicing-by-8) when it has been built."""
    global _native
    if _native is None:
        _native = False
        try:
            import ctypes
            from .. import ops
            if ops.available():
                fn = ops.lib().dtg_crc32c
                fn.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_uint32]
                fn.restype = ctypes.c_uint32
                _native = fn
        except Exception:                                  # noqa: BLE001 - pure-Python fallback
            _native = False
    return _native


def crc32c(data: bytes, crc: int = 0) -> int:
    """CRC-32C (Castagnoli), the checksum leveldb and TF use."""
    fn = _native_crc() if len(data) >= 256 else None
    if fn:
        return int(fn(bytes(data), len(data), crc))
    c = crc ^ 0xFFFFFFFF
    t = _TABLE
    for b in data:
        c = t[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc32c(data: bytes) -> int:
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xa282ead8) & 0xFFFFFFFF


def _unmask(m: int) -> int:
    rot = (m - 0xa282ead8) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


# ----------------------------------------------------------------- varint / proto
def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf: bytes, pos: int) -> Tuple[int, int]:
    shift = val = 0
    while True:
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7


def _field_varint(num: int, val: int) -> bytes:
    return _varint(num << 3) + _varint(val)


def _field_bytes(num: int, val: bytes) -> bytes:
    return _varint((num << 3) | 2) + _varint(len(val)) + val


def _parse_proto(buf: bytes) -> Dict[int, list]:
    """Minimal protobuf wire parser: {field: [values]} (varint -> int, len-delimited -> bytes,
    fixed32 -> int)."""
    out: Dict[int, list] = {}
    pos = 0
    while pos < len(buf):
        tag, pos = _read_varint(buf, pos)
        num, wt = tag >> 3, tag & 7
        if wt == 0:
            v, pos = _read_varint(buf, pos)
        elif wt == 2:
            n, pos = _read_varint(buf, pos)
            v = buf[pos:pos + n]
            pos += n
        elif wt == 5:
            v = struct.unpack_from("<I", buf, pos)[0]
            pos += 4
        elif wt == 1:
            v = struct.unpack_from("<Q", buf, pos)[0]
            pos += 8
        else:
            raise ValueError("unsupported wire type %d" % wt)
        out.setdefault(num, []).append(v)
    return out


def _header_proto() -> bytes:
    # BundleHeaderProto: num_shards = 1, endianness = LITTLE (0, default: omitted),
    # version = VersionDef{producer = 1}
    return _field_varint(1, 1) + _field_bytes(3, _field_varint(1, 1))


def _entry_proto(dtype: int, shape, offset: int, size: int, crc: int) -> bytes:
    dims = b"".join(_field_bytes(2, _field_varint(1, int(d))) for d in shape)   # TensorShapeProto.dim
    out = _field_varint(1, dtype) + _field_bytes(2, dims)
    # shard_id = 0 (default: omitted)
    if offset:
        out += _field_varint(4, offset)
    out += _field_varint(5, size)
    out += _varint((6 << 3) | 5) + struct.pack("<I", crc)                        # fixed32 crc32c
    return out


# ------------------------------------------------------------------- table writer
class _BlockBuilder:
    def __init__(self):
        self.buf = bytearray()
        self.restarts = [0]
        self.count = 0
        self.last_key = b""

    def add(self, key: bytes, value: bytes) -> None:
        shared = 0
        if self.count % _RESTART == 0 and self.count:
            self.restarts.append(len(self.buf))
        elif self.count:
            lim = min(len(key), len(self.last_key))
            while shared < lim and key[shared] == self.last_key[shared]:
                shared += 1
        self.buf += _varint(shared) + _varint(len(key) - shared) + _varint(len(value))
        self.buf += key[shared:] + value
        self.last_key = key
        self.count += 1

    def size(self) -> int:
        return len(self.buf) + 4 * len(self.restarts) + 4

    def finish(self) -> bytes:
        out = bytes(self.buf) + b"".join(struct.pack("<I", r) for r in self.restarts)
        return out + struct.pack("<I", len(self.restarts))


def _write_block(f, block: bytes) -> Tuple[int, int]:
    off = f.tell()
    f.write(block)
    f.write(b"\x00")                                                # no compression
    f.write(struct.pack("<I", masked_crc32c(block + b"\x00")))
    return off, len(block)


def _handle(off: int, size: int) -> bytes:
    return _varint(off) + _varint(size)


def _write_table(path: str, items: List[Tuple[bytes, bytes]]) -> None:
    """``items`` sorted by key."""
    with open(path, "wb") as f:
        index = _BlockBuilder()
        blk = _BlockBuilder()
        for key, value in items:
            blk.add(key, value)
            if blk.size() >= _BLOCK_SIZE:
                off, size = _write_block(f, blk.finish())
                index.add(blk.last_key, _handle(off, size))
                blk = _BlockBuilder()
        if blk.count or not items:
            off, size = _write_block(f, blk.finish())
            index.add(blk.last_key, _handle(off, size))
        meta_off, meta_size = _write_block(f, _BlockBuilder().finish())
        idx_off, idx_size = _write_block(f, index.finish())
        footer = _handle(meta_off, meta_size) + _handle(idx_off, idx_size)
        footer += b"\x00" * (40 - len(footer))
        footer += struct.pack("<Q", _MAGIC)
        f.write(footer)
        f.flush()
        os.fsync(f.fileno())


# ------------------------------------------------------------------- table reader
def _read_block(buf: bytes, off: int, size: int, verify: bool = True) -> bytes:
    block = buf[off:off + size]
    ctype = buf[off + size]
    (crc,) = struct.unpack_from("<I", buf, off + size + 1)
    if ctype != 0:
        raise ValueError("compressed table blocks are not supported (type %d)" % ctype)
    if verify and _unmask(crc) != crc32c(block + bytes([ctype])):
        raise ValueError("table block checksum mismatch at offset %d" % off)
    return block


def _block_entries(block: bytes) -> List[Tuple[bytes, bytes]]:
    (nrestarts,) = struct.unpack_from("<I", block, len(block) - 4)
    end = len(block) - 4 - 4 * nrestarts
    out, pos, key = [], 0, b""
    while pos < end:
        shared, pos = _read_varint(block, pos)
        non_shared, pos = _read_varint(block, pos)
        vlen, pos = _read_varint(block, pos)
        key = key[:shared] + block[pos:pos + non_shared]
        pos += non_shared
        out.append((key, block[pos:pos + vlen]))
        pos += vlen
    return out


def _read_table(path: str) -> List[Tuple[bytes, bytes]]:
    with open(path, "rb") as f:
        buf = f.read()
    if len(buf) < 48 or struct.unpack_from("<Q", buf, len(buf) - 8)[0] != _MAGIC:
        raise ValueError("%s is not a TF checkpoint index (bad table magic)" % path)
    footer = buf[-48:]
    _, pos = _read_varint(footer, 0)
    _, pos = _read_varint(footer, pos)
    idx_off, pos = _read_varint(footer, pos)
    idx_size, pos = _read_varint(footer, pos)
    items = []
    for _, handle in _block_entries(_read_block(buf, idx_off, idx_size)):
        off, p = _read_varint(handle, 0)
        size, _ = _read_varint(handle, p)
        items.extend(_block_entries(_read_block(buf, off, size)))
    return items


# ------------------------------------------------------------------------ public
def _data_path(prefix: str) -> str:
    return prefix + ".data-00000-of-00001"


def is_bundle(prefix: str) -> bool:
    return os.path.exists(prefix + ".index") and os.path.exists(_data_path(prefix))


def write_bundle(prefix: str, tensors: Dict[str, torch.Tensor]) -> None:
    """Write ``{variable name: tensor}`` as ``<prefix>.index`` + ``<prefix>.data-00000-of-00001``."""
    items: List[Tuple[bytes, bytes]] = [(b"", _header_proto())]
    tmp_data = _data_path(prefix) + ".tmp.%d" % os.getpid()
    off = 0
    with open(tmp_data, "wb") as f:
        for name in sorted(tensors, key=lambda s: s.encode()):
            t = tensors[name].detach().cpu().contiguous()
            if t.dtype not in _DT:
                raise TypeError("dtype %s has no TF DataType mapping" % t.dtype)
            if t.dtype == torch.bfloat16:
                raw = t.view(torch.int16).numpy().astype("<i2").tobytes()
            else:
                raw = t.numpy().astype(_NP[t.dtype]).tobytes()
            f.write(raw)
            items.append((name.encode(), _entry_proto(_DT[t.dtype], t.shape, off, len(raw),
                                                      masked_crc32c(raw))))
            off += len(raw)
        f.flush()
        os.fsync(f.fileno())
    tmp_index = prefix + ".index.tmp.%d" % os.getpid()
    _write_table(tmp_index, items)
    os.replace(tmp_data, _data_path(prefix))
    os.replace(tmp_index, prefix + ".index")


def _entries(prefix: str) -> Dict[str, Dict[int, list]]:
    items = _read_table(prefix + ".index")
    if not items or items[0][0] != b"":
        raise ValueError("bundle header entry missing in %s.index" % prefix)
    header = _parse_proto(items[0][1])
    if header.get(1, [1])[0] != 1:
        raise ValueError("multi-shard bundles are not supported (num_shards=%d)" % header[1][0])
    if header.get(2, [0])[0] != 0:
        raise ValueError("big-endian bundles are not supported")
    return {k.decode(): _parse_proto(v) for k, v in items[1:]}


def _shape_of(entry: Dict[int, list]) -> Tuple[int, ...]:
    if 2 not in entry:
        return ()
    dims = _parse_proto(entry[2][0]).get(2, [])
    return tuple(_parse_proto(d).get(1, [0])[0] for d in dims)


def list_bundle(prefix: str) -> Dict[str, Tuple[torch.dtype, Tuple[int, ...]]]:
    return {k: (_DT_INV[e[1][0]], _shape_of(e)) for k, e in _entries(prefix).items()}


def read_bundle(prefix: str, names: Optional[Iterable[str]] = None,
                verify: bool = True) -> Dict[str, torch.Tensor]:
    """Read tensors (all, or ``names``) from a TF checkpoint V2 prefix."""
    want = set(names) if names is not None else None
    out: Dict[str, torch.Tensor] = {}
    with open(_data_path(prefix), "rb") as f:
        for name, e in _entries(prefix).items():
            if want is not None and name not in want:
                continue
            if 7 in e:
                raise ValueError("%s is a partitioned (sliced) variable: not supported" % name)
            dt = _DT_INV.get(e[1][0])
            if dt is None:
                raise TypeError("%s: TF DataType %d is not supported" % (name, e[1][0]))
            shape = _shape_of(e)
            off, size = e.get(4, [0])[0], e.get(5, [0])[0]
            f.seek(off)
            raw = f.read(size)
            if verify and 6 in e and _unmask(e[6][0]) != crc32c(raw):
                raise ValueError("%s: tensor checksum mismatch" % name)
            if dt == torch.bfloat16:
                a = np.frombuffer(raw, dtype="<i2").reshape(shape).copy()
                out[name] = torch.from_numpy(a).view(torch.bfloat16)
            else:
                a = np.frombuffer(raw, dtype=_NP[dt]).reshape(shape).copy()
                out[name] = torch.from_numpy(a)
    return out
