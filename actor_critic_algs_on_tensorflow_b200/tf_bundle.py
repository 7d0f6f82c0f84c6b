"""TensorFlow V2 checkpoint ("tensor bundle") reader and writer, no TensorFlow needed.

The reference saves and restores with ``tf.train.Saver`` (``A3C/process.py:140-144,197,
211-214``; ``Basic_AC/run_AC.py:199,282-283``); the shipped fixture is
``demos/model-Pendulum_a3c.{index,data-00000-of-00001}``.  This module reads and writes
exactly that on-disk format, so ``test_model.py`` semantics keep
working: a checkpoint written here is byte-compatible with one written by TF 1.3 for the
same tensors (``write_bundle(read_bundle(demo))`` reproduces the demo files byte for byte --
tests/test_tf_bundle.py).

Format: ``<prefix>.index`` is a leveldb-style SSTable (prefix-compressed keys, restart
interval 16, 5-byte block trailers with masked crc32c, 48-byte footer); key "" holds a
``BundleHeaderProto``, every other key a ``BundleEntryProto`` (dtype, shape, offset, size,
masked crc32c of the raw bytes).  ``<prefix>.data-00000-of-00001`` is the raw little-endian
tensors concatenated in byte-sorted key order.
"""
import os
import struct

import numpy as np

_MAGIC = 0xDB4775248B80FB57
_RESTART_INTERVAL = 16
_DT_FLOAT, _DT_INT32, _DT_INT64 = 1, 3, 9
_DTYPES = {_DT_FLOAT: np.dtype("<f4"), _DT_INT32: np.dtype("<i4"), _DT_INT64: np.dtype("<i8"), 2: np.dtype("<f8")}
_DT_OF = {np.dtype("float32"): _DT_FLOAT, np.dtype("int32"): _DT_INT32, np.dtype("int64"): _DT_INT64,
          np.dtype("float64"): 2}


# ------------------------------------------------------------------ crc32c (Castagnoli), masked as in leveldb/TF
def _make_table():
    tbl = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
        tbl.append(c)
    return np.array(tbl, dtype=np.uint32)


_TABLE = _make_table()
_TABLE_L = [int(x) for x in _TABLE]


def crc32c(data: bytes) -> int:
    c = 0xFFFFFFFF
    tbl = _TABLE_L
    for b in data:
        c = tbl[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc32c(data: bytes) -> int:
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


# ------------------------------------------------------------------ varints / minimal protobuf
def _put_varint(v: int) -> bytes:
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _get_varint(buf, pos):
    shift = v = 0
    while True:
        b = buf[pos]
        pos += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, pos
        shift += 7


def _parse_proto(buf):
    """-> list of (field, wire_type, value) with value int (varint/fixed) or bytes."""
    pos, out = 0, []
    while pos < len(buf):
        tag, pos = _get_varint(buf, pos)
        field, wt = tag >> 3, tag & 7
        if wt == 0:
            v, pos = _get_varint(buf, pos)
        elif wt == 2:
            n, pos = _get_varint(buf, pos)
            v = bytes(buf[pos:pos + n])
            pos += n
        elif wt == 5:
            v = struct.unpack_from("<I", buf, pos)[0]
            pos += 4
        elif wt == 1:
            v = struct.unpack_from("<Q", buf, pos)[0]
            pos += 8
        else:
            raise ValueError("unsupported protobuf wire type %d" % wt)
        out.append((field, wt, v))
    return out


def _entry_proto(dtype_enum, shape, offset, size, crc):
    shp = b"".join(b"\x12" + _put_varint(len(d)) + d for d in (b"\x08" + _put_varint(s) for s in shape))
    out = b"\x08" + _put_varint(dtype_enum) + b"\x12" + _put_varint(len(shp)) + shp
    if offset:
        out += b"\x20" + _put_varint(offset)
    out += b"\x28" + _put_varint(size) + b"\x35" + struct.pack("<I", crc)
    return out


_HEADER_PROTO = b"\x08\x01\x1a\x02\x08\x01"     # num_shards=1, version{producer=1}


# ------------------------------------------------------------------ SSTable blocks
def _build_block(items):
    out, restarts, last = bytearray(), [], b""
    for i, (k, v) in enumerate(items):
        if i % _RESTART_INTERVAL == 0:
            restarts.append(len(out))
            shared = 0
        else:
            shared = 0
            while shared < min(len(k), len(last)) and k[shared] == last[shared]:
                shared += 1
        out += _put_varint(shared) + _put_varint(len(k) - shared) + _put_varint(len(v)) + k[shared:] + v
        last = k
    if not restarts:
        restarts = [0]
    for r in restarts:
        out += struct.pack("<I", r)
    out += struct.pack("<I", len(restarts))
    return bytes(out)


def _with_trailer(block):
    return block + b"\x00" + struct.pack("<I", masked_crc32c(block + b"\x00"))


def _parse_block(buf):
    n_restarts = struct.unpack_from("<I", buf, len(buf) - 4)[0]
    end = len(buf) - 4 - 4 * n_restarts
    pos, key, items = 0, b"", []
    while pos < end:
        shared, pos = _get_varint(buf, pos)
        non_shared, pos = _get_varint(buf, pos)
        vlen, pos = _get_varint(buf, pos)
        key = key[:shared] + bytes(buf[pos:pos + non_shared])
        pos += non_shared
        items.append((key, bytes(buf[pos:pos + vlen])))
        pos += vlen
    return items


def _read_block(data, offset, size, verify=True):
    blk = data[offset:offset + size]
    typ = data[offset + size]
    crc = struct.unpack_from("<I", data, offset + size + 1)[0]
    if typ != 0:
        raise ValueError("compressed SSTable blocks are not supported")
    if verify and masked_crc32c(blk + bytes([typ])) != crc:
        raise ValueError("SSTable block checksum mismatch")
    return blk


def _short_successor(key: bytes) -> bytes:
    """leveldb BytewiseComparator::FindShortSuccessor."""
    for i, b in enumerate(key):
        if b != 0xFF:
            return key[:i] + bytes([b + 1])
    return key


# ------------------------------------------------------------------ public API
def read_bundle(prefix, verify=True):
    """Returns ``{tensor_name: np.ndarray}`` (insertion order = byte-sorted key order).
    Raises FileNotFoundError / ValueError (bad magic, checksum mismatch) like a failed
    ``saver.restore``."""
    with open(prefix + ".index", "rb") as fh:
        idx = fh.read()
    if len(idx) < 48 or struct.unpack_from("<Q", idx, len(idx) - 8)[0] != _MAGIC:
        raise ValueError("%s.index: not a TF V2 checkpoint index (bad magic)" % prefix)
    foot = idx[-48:]
    _, p = _get_varint(foot, 0)
    _, p = _get_varint(foot, p)
    ioff, p = _get_varint(foot, p)
    isize, p = _get_varint(foot, p)
    entries = []
    for _, handle in _parse_block(_read_block(idx, ioff, isize, verify)):
        boff, q = _get_varint(handle, 0)
        bsize, q = _get_varint(handle, q)
        entries += _parse_block(_read_block(idx, boff, bsize, verify))
    shards, out = {}, {}
    num_shards = 1
    for key, val in entries:
        if key == b"":
            for f, _, v in _parse_proto(val):
                if f == 1:
                    num_shards = v
                if f == 2 and v != 0:
                    raise ValueError("big-endian bundles are not supported")
            continue
        dtype_enum, shape, shard, offset, size, crc = 0, [], 0, 0, 0, None
        for f, _, v in _parse_proto(val):
            if f == 1:
                dtype_enum = v
            elif f == 2:
                for f2, _, d in _parse_proto(v):
                    if f2 == 2:
                        sz = [x for g, _, x in _parse_proto(d) if g == 1]
                        shape.append(sz[0] if sz else 0)
            elif f == 3:
                shard = v
            elif f == 4:
                offset = v
            elif f == 5:
                size = v
            elif f == 6:
                crc = v
        if dtype_enum not in _DTYPES:
            raise ValueError("tensor %r: unsupported dtype enum %d" % (key, dtype_enum))
        if shard not in shards:
            with open("%s.data-%05d-of-%05d" % (prefix, shard, num_shards), "rb") as fh:
                shards[shard] = fh.read()
        raw = shards[shard][offset:offset + size]
        if len(raw) != size:
            raise ValueError("tensor %r: data file truncated" % key)
        if verify and crc is not None and masked_crc32c(raw) != crc:
            raise ValueError("tensor %r: crc32c mismatch" % key)
        out[key.decode()] = np.frombuffer(raw, dtype=_DTYPES[dtype_enum]).reshape(shape).copy()
    return out


def write_bundle(prefix, tensors):
    """Write ``{name: array}`` as ``<prefix>.index`` + ``<prefix>.data-00000-of-00001``."""
    d = os.path.dirname(prefix)
    if d:
        os.makedirs(d, exist_ok=True)
    items, data, offset = [(b"", _HEADER_PROTO)], bytearray(), 0
    for name in sorted(tensors, key=lambda s: s.encode()):
        arr = np.asarray(tensors[name])
        if arr.dtype not in _DT_OF:
            arr = arr.astype(np.float32)
        raw = np.ascontiguousarray(arr.astype(arr.dtype.newbyteorder("<"))).tobytes()
        items.append((name.encode(), _entry_proto(_DT_OF[arr.dtype], arr.shape, offset, len(raw), masked_crc32c(raw))))
        data += raw
        offset += len(raw)
    block = _build_block(items)
    out = bytearray(_with_trailer(block))
    meta_off = len(out)
    meta = _build_block([])
    out += _with_trailer(meta)
    index_off = len(out)
    handle = _put_varint(0) + _put_varint(len(block))
    index = _build_block([(_short_successor(items[-1][0]), handle)])
    out += _with_trailer(index)
    foot = _put_varint(meta_off) + _put_varint(len(meta)) + _put_varint(index_off) + _put_varint(len(index))
    out += foot + b"\x00" * (40 - len(foot)) + struct.pack("<Q", _MAGIC)
    with open(prefix + ".data-00000-of-00001", "wb") as fh:
        fh.write(bytes(data))
    with open(prefix + ".index", "wb") as fh:
        fh.write(bytes(out))


def write_checkpoint_state(directory, latest, all_paths):
    """The ``checkpoint`` text file ``tf.train.Saver`` maintains next to the bundles."""
    with open(os.path.join(directory, "checkpoint"), "w") as fh:
        fh.write('model_checkpoint_path: "%s"\n' % latest)
        for p in all_paths:
            fh.write('all_model_checkpoint_paths: "%s"\n' % p)
