"""Standalone reader/writer for TensorFlow's V2 checkpoint ("tensor bundle") format, plus the
Estimator-style directory management around it.

The reference never touches checkpoints directly -- ``tf.estimator.Estimator(model_dir=...)``
does (HVD.py:403-415, PS.py:492-497): ``model.ckpt-<step>.index`` + ``.data-00000-of-00001``,
a text-proto ``checkpoint`` state file, keep the last 5, restore the newest on start.  This
module writes and reads those same files without TensorFlow, keyed by the reference's
variable names, so a checkpoint of this framework can be opened by
``tf.train.load_checkpoint`` and vice versa.

Format (tensorflow/core/util/tensor_bundle):
  <prefix>.index  a LevelDB-format SSTable: key "" -> BundleHeaderProto, key <name> ->
                  BundleEntryProto{dtype, shape, shard_id, offset, size, crc32c}
  <prefix>.data-00000-of-00001  raw little-endian tensor bytes, back to back.
"""
from __future__ import annotations

import os
import re
import struct
import time
from collections import OrderedDict
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np

from ..data.tfrecord import crc32c as _crc32c

TABLE_MAGIC = 0xDB4775248B80FB57
_BLOCK_SIZE = 256 * 1024
_RESTART_INTERVAL = 16

# tensorflow DataType enum values
_DT = {np.dtype("float32"): 1, np.dtype("float64"): 2, np.dtype("int32"): 3, np.dtype("int64"): 9,
       np.dtype("uint8"): 4, np.dtype("int16"): 5, np.dtype("int8"): 6, np.dtype("bool"): 10,
       np.dtype("float16"): 19}
_DT_INV = {v: k for k, v in _DT.items()}


def _mask(crc: int) -> int:
    return (((crc >> 15) | (crc << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def _unmask(m: int) -> int:
    rot = (m - 0xA282EAD8) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


def _crc_extend(data: bytes) -> int:
    return _crc32c(data)


# ------------------------------------------------------------------ protobuf bits
def _varint(v: int) -> bytes:
    out = bytearray()
    v &= (1 << 64) - 1
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _read_varint(buf: bytes, pos: int) -> Tuple[int, int]:
    r, shift = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        r |= (b & 0x7F) << shift
        if not b & 0x80:
            return r, pos
        shift += 7


def _pb_fields(buf: bytes) -> Iterable[Tuple[int, int, object]]:
    pos = 0
    while pos < len(buf):
        tag, pos = _read_varint(buf, pos)
        field, wire = tag >> 3, tag & 7
        if wire == 0:
            v, pos = _read_varint(buf, pos)
        elif wire == 1:
            v = buf[pos:pos + 8]
            pos += 8
        elif wire == 2:
            n, pos = _read_varint(buf, pos)
            v = buf[pos:pos + n]
            pos += n
        elif wire == 5:
            v = buf[pos:pos + 4]
            pos += 4
        else:
            raise ValueError(f"unsupported protobuf wire type {wire}")
        yield field, wire, v


def _encode_shape(shape: Tuple[int, ...]) -> bytes:
    out = bytearray()
    for d in shape:
        dim = b"\x08" + _varint(d)                 # Dim.size = 1
        out += b"\x12" + _varint(len(dim)) + dim   # TensorShapeProto.dim = 2
    return bytes(out)


def _decode_shape(buf: bytes) -> Tuple[int, ...]:
    dims = []
    for f, _, v in _pb_fields(buf):
        if f == 2:
            size = 0
            for f2, _, v2 in _pb_fields(v):
                if f2 == 1:
                    size = v2
            dims.append(size)
    return tuple(dims)


def _encode_entry(dtype: np.dtype, shape, offset: int, size: int, crc: int) -> bytes:
    sh = _encode_shape(tuple(shape))
    out = bytearray()
    out += b"\x08" + _varint(_DT[np.dtype(dtype)])
    out += b"\x12" + _varint(len(sh)) + sh
    # shard_id = 0 is the default and is omitted
    if offset:
        out += b"\x20" + _varint(offset)
    out += b"\x28" + _varint(size)
    out += b"\x35" + struct.pack("<I", crc)
    return bytes(out)


def _encode_header() -> bytes:
    # num_shards = 1, endianness LITTLE (0, omitted), version {producer: 1}
    return b"\x08\x01" + b"\x1a\x02\x08\x01"


# ------------------------------------------------------------------ SSTable
class _BlockBuilder:
    def __init__(self):
        self.buf = bytearray()
        self.restarts = [0]
        self.count = 0
        self.last_key = b""

    def add(self, key: bytes, value: bytes) -> None:
        shared = 0
        if self.count % _RESTART_INTERVAL == 0 and self.count:
            self.restarts.append(len(self.buf))
        elif self.count:
            m = min(len(key), len(self.last_key))
            while shared < m and key[shared] == self.last_key[shared]:
                shared += 1
        self.buf += _varint(shared) + _varint(len(key) - shared) + _varint(len(value))
        self.buf += key[shared:] + value
        self.last_key = key
        self.count += 1

    def finish(self) -> bytes:
        out = bytes(self.buf)
        for r in self.restarts:
            out += struct.pack("<I", r)
        return out + struct.pack("<I", len(self.restarts))

    def size_estimate(self) -> int:
        return len(self.buf) + 4 * len(self.restarts) + 4


def _write_block(f, contents: bytes) -> Tuple[int, int]:
    offset = f.tell()
    trailer_type = b"\x00"
    crc = _mask(_crc_extend(contents + trailer_type))
    f.write(contents + trailer_type + struct.pack("<I", crc))
    return offset, len(contents)


def _write_table(path: str, items: List[Tuple[bytes, bytes]]) -> None:
    with open(path, "wb") as f:
        index = _BlockBuilder()
        blk = _BlockBuilder()
        for k, v in items:
            blk.add(k, v)
            if blk.size_estimate() >= _BLOCK_SIZE:
                off, sz = _write_block(f, blk.finish())
                index.add(blk.last_key, _varint(off) + _varint(sz))
                blk = _BlockBuilder()
        if blk.count:
            off, sz = _write_block(f, blk.finish())
            index.add(blk.last_key, _varint(off) + _varint(sz))
        meta_off, meta_sz = _write_block(f, _BlockBuilder().finish())
        idx_off, idx_sz = _write_block(f, index.finish())
        footer = _varint(meta_off) + _varint(meta_sz) + _varint(idx_off) + _varint(idx_sz)
        footer += b"\x00" * (40 - len(footer))
        f.write(footer + struct.pack("<Q", TABLE_MAGIC))


def _parse_block(buf: bytes) -> List[Tuple[bytes, bytes]]:
    n_restarts = struct.unpack_from("<I", buf, len(buf) - 4)[0]
    end = len(buf) - 4 - 4 * n_restarts
    pos, key, out = 0, b"", []
    while pos < end:
        shared, pos = _read_varint(buf, pos)
        non_shared, pos = _read_varint(buf, pos)
        vlen, pos = _read_varint(buf, pos)
        key = key[:shared] + buf[pos:pos + non_shared]
        pos += non_shared
        out.append((key, buf[pos:pos + vlen]))
        pos += vlen
    return out


def _read_block(data: bytes, off: int, size: int, verify: bool = True) -> bytes:
    contents = data[off:off + size]
    if verify:
        typ = data[off + size:off + size + 1]
        crc = struct.unpack_from("<I", data, off + size + 1)[0]
        if typ != b"\x00":
            raise ValueError("compressed SSTable blocks are not supported")
        if _mask(_crc_extend(contents + typ)) != crc:
            raise ValueError("checkpoint index block CRC mismatch")
    return contents


def _read_table(path: str) -> List[Tuple[bytes, bytes]]:
    data = open(path, "rb").read()
    if len(data) < 48 or struct.unpack_from("<Q", data, len(data) - 8)[0] != TABLE_MAGIC:
        raise ValueError(f"{path} is not a tensor-bundle index (bad magic)")
    footer = data[-48:-8]
    pos = 0
    _, pos = _read_varint(footer, pos)
    _, pos = _read_varint(footer, pos)
    idx_off, pos = _read_varint(footer, pos)
    idx_sz, pos = _read_varint(footer, pos)
    items = []
    for _, handle in _parse_block(_read_block(data, idx_off, idx_sz)):
        boff, p = _read_varint(handle, 0)
        bsz, _ = _read_varint(handle, p)
        items.extend(_parse_block(_read_block(data, boff, bsz)))
    return items


# ------------------------------------------------------------------ bundle API
def write_bundle(prefix: str, tensors: Dict[str, np.ndarray]) -> None:
    """Writes ``<prefix>.index`` and ``<prefix>.data-00000-of-00001`` atomically (tmp + rename)."""
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    data_path = f"{prefix}.data-00000-of-00001"
    items: List[Tuple[bytes, bytes]] = [(b"", _encode_header())]
    offset = 0
    with open(data_path + ".tmp", "wb") as f:
        for name in sorted(tensors.keys(), key=lambda s: s.encode()):
            a = np.asarray(tensors[name])
            if not a.flags.c_contiguous:
                a = np.ascontiguousarray(a)
            if a.dtype not in _DT:
                raise TypeError(f"{name}: dtype {a.dtype} is not supported in checkpoints")
            raw = a.tobytes()
            f.write(raw)
            items.append((name.encode(), _encode_entry(a.dtype, a.shape, offset, len(raw), _mask(_crc32c(raw)))))
            offset += len(raw)
    _write_table(prefix + ".index.tmp", items)
    os.replace(data_path + ".tmp", data_path)
    os.replace(prefix + ".index.tmp", prefix + ".index")


def list_bundle(prefix: str) -> "OrderedDict[str, Tuple[np.dtype, Tuple[int, ...], int, int, int]]":
    """name -> (dtype, shape, offset, size, masked_crc)."""
    out: "OrderedDict[str, tuple]" = OrderedDict()
    for k, v in _read_table(prefix + ".index"):
        if k == b"":
            continue
        dtype, shape, off, size, crc = np.dtype("float32"), (), 0, 0, 0
        for f, _, val in _pb_fields(v):
            if f == 1:
                dtype = _DT_INV[val]
            elif f == 2:
                shape = _decode_shape(val)
            elif f == 4:
                off = val
            elif f == 5:
                size = val
            elif f == 6:
                crc = struct.unpack("<I", val)[0]
        out[k.decode()] = (dtype, shape, off, size, crc)
    return out


def read_bundle(prefix: str, names: Optional[Iterable[str]] = None, verify_crc: bool = True) -> Dict[str, np.ndarray]:
    entries = list_bundle(prefix)
    want = list(entries.keys()) if names is None else list(names)
    out: Dict[str, np.ndarray] = {}
    with open(f"{prefix}.data-00000-of-00001", "rb") as f:
        for n in want:
            if n not in entries:
                raise KeyError(f"{n} not found in checkpoint {prefix}")
            dtype, shape, off, size, crc = entries[n]
            f.seek(off)
            raw = f.read(size)
            if len(raw) != size:
                raise ValueError(f"{prefix}: data file truncated while reading {n}")
            if verify_crc and _mask(_crc32c(raw)) != crc:
                raise ValueError(f"{prefix}: CRC mismatch for tensor {n}")
            out[n] = np.frombuffer(raw, dtype=dtype).reshape(shape).copy()
    return out


# ------------------------------------------------------------------ directory management
_STATE_FILE = "checkpoint"


def _read_state(model_dir: str) -> Tuple[Optional[str], List[str]]:
    path = os.path.join(model_dir, _STATE_FILE)
    if not os.path.exists(path):
        return None, []
    latest, allp = None, []
    for line in open(path):
        m = re.match(r'\s*(model_checkpoint_path|all_model_checkpoint_paths)\s*:\s*"(.*)"', line)
        if m:
            if m.group(1) == "model_checkpoint_path":
                latest = m.group(2)
            else:
                allp.append(m.group(2))
    return latest, allp


def latest_checkpoint(model_dir: str) -> Optional[str]:
    """Prefix of the newest checkpoint (``tf.train.latest_checkpoint``), or None."""
    if not model_dir or not os.path.isdir(model_dir):
        return None
    latest, _ = _read_state(model_dir)
    if latest is None:
        return None
    prefix = latest if os.path.isabs(latest) else os.path.join(model_dir, latest)
    return prefix if os.path.exists(prefix + ".index") else None


def checkpoint_step(prefix: str) -> int:
    m = re.search(r"-(\d+)$", prefix)
    return int(m.group(1)) if m else 0


class CheckpointManager:
    """Estimator-like saver: ``model.ckpt-<global_step>``, state file, keep_max, time-based period."""

    def __init__(self, model_dir: str, keep_max: int = 5, save_secs: int = 600):
        self.model_dir = model_dir
        self.keep_max = keep_max
        self.save_secs = save_secs
        self._last_save = time.time()
        if model_dir:
            os.makedirs(model_dir, exist_ok=True)

    def due(self) -> bool:
        return bool(self.model_dir) and self.save_secs > 0 and time.time() - self._last_save >= self.save_secs

    def save(self, step: int, tensors: Dict[str, np.ndarray]) -> str:
        name = f"model.ckpt-{int(step)}"
        prefix = os.path.join(self.model_dir, name)
        t = dict(tensors)
        t["global_step"] = np.asarray(int(step), dtype=np.int64)
        write_bundle(prefix, t)
        _, allp = _read_state(self.model_dir)
        allp = [p for p in allp if p != name] + [name]
        drop, allp = (allp[:-self.keep_max], allp[-self.keep_max:]) if self.keep_max > 0 else ([], allp)
        tmp = os.path.join(self.model_dir, _STATE_FILE + ".tmp")
        with open(tmp, "w") as f:
            f.write(f'model_checkpoint_path: "{name}"\n')
            for p in allp:
                f.write(f'all_model_checkpoint_paths: "{p}"\n')
        os.replace(tmp, os.path.join(self.model_dir, _STATE_FILE))
        for p in drop:
            for suffix in (".index", ".data-00000-of-00001"):
                try:
                    os.remove(os.path.join(self.model_dir, p + suffix))
                except OSError:
                    pass
        self._last_save = time.time()
        return prefix

    def restore_latest(self) -> Optional[Tuple[int, Dict[str, np.ndarray]]]:
        prefix = latest_checkpoint(self.model_dir)
        if prefix is None:
            return None
        t = read_bundle(prefix)
        step = int(t.pop("global_step")) if "global_step" in t else checkpoint_step(prefix)
        return step, t
