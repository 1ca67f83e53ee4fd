"""TFRecord + tf.train.Example access through libdfm_io.so (no TensorFlow).

Python face of the native reader/writer/codec that replaces
``tf.data.TFRecordDataset`` (reference HVD.py:128), ``tf.parse_example``
(HVD.py:109-118) and ``tf.python_io.TFRecordWriter`` (tools/libsvm_to_tfrecord.py:29).
On-disk keys are ``label`` / ``ids`` / ``values``; the in-memory names
``feat_ids`` / ``feat_vals`` exist only on the model side (HVD.py:124).
"""
from __future__ import annotations

import ctypes as C
from typing import Iterator, Optional, Tuple

import numpy as np

from .. import _native


class DataError(RuntimeError):
    pass


def _lib():
    lib = _native.io()
    if not getattr(lib, "_tfrecord_typed", False):
        lib.dfio_crc32c.restype = C.c_uint32
        lib.dfio_crc32c.argtypes = [C.c_void_p, C.c_size_t]
        lib.dfio_masked_crc32c.restype = C.c_uint32
        lib.dfio_masked_crc32c.argtypes = [C.c_void_p, C.c_size_t]
        lib.dfio_reader_open.restype = C.c_void_p
        lib.dfio_reader_open.argtypes = [C.c_char_p, C.c_int]
        lib.dfio_reader_next.argtypes = [C.c_void_p, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_uint64)]
        lib.dfio_reader_close.argtypes = [C.c_void_p]
        lib.dfio_count_records.restype = C.c_int64
        lib.dfio_count_records.argtypes = [C.c_char_p, C.c_int]
        lib.dfio_writer_open.restype = C.c_void_p
        lib.dfio_writer_open.argtypes = [C.c_char_p]
        lib.dfio_writer_write.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        lib.dfio_writer_write_example.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_int]
        lib.dfio_writer_close.argtypes = [C.c_void_p]
        lib.dfio_encode_example.restype = C.c_int64
        lib.dfio_encode_example.argtypes = [C.c_float, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_uint64, C.c_int]
        lib.dfio_parse_example.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        lib.dfio_read_all.restype = C.c_int64
        lib.dfio_read_all.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_int,
                                      C.c_void_p, C.c_void_p, C.c_int]
        lib.dfio_libsvm_to_tfrecord.restype = C.c_int64
        lib.dfio_libsvm_to_tfrecord.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        lib.dfio_generate_synthetic.restype = C.c_int64
        lib.dfio_generate_synthetic.argtypes = [C.c_char_p, C.c_int64, C.c_int, C.c_int64, C.c_int, C.c_uint64,
                                                C.c_double, C.c_double]
        lib._tfrecord_typed = True
    return lib


def _err() -> str:
    return _lib().dfio_last_error().decode("utf-8", "replace")


def crc32c(data: bytes) -> int:
    return int(_lib().dfio_crc32c(data, len(data)))


def masked_crc32c(data: bytes) -> int:
    return int(_lib().dfio_masked_crc32c(data, len(data)))


def count_records(path: str, verify_crc: bool = True) -> int:
    n = _lib().dfio_count_records(path.encode(), int(verify_crc))
    if n < 0:
        raise DataError(_err())
    return int(n)


def iter_records(path: str, verify_crc: bool = True) -> Iterator[bytes]:
    """Yields raw record payloads (serialized Examples).  ``path`` may be a FIFO or "-"."""
    lib = _lib()
    h = lib.dfio_reader_open(path.encode(), int(verify_crc))
    if not h:
        raise DataError(_err())
    try:
        p = C.POINTER(C.c_uint8)()
        n = C.c_uint64()
        while True:
            r = lib.dfio_reader_next(h, C.byref(p), C.byref(n))
            if r == 0:
                return
            if r < 0:
                raise DataError(_err())
            yield C.string_at(p, n.value)
    finally:
        lib.dfio_reader_close(h)


def encode_example(label: float, ids, vals, packed: bool = True) -> bytes:
    ids = np.ascontiguousarray(ids, dtype=np.int64)
    vals = np.ascontiguousarray(vals, dtype=np.float32)
    assert ids.shape == vals.shape and ids.ndim == 1
    cap = 64 + ids.size * 16
    buf = C.create_string_buffer(cap)
    n = _lib().dfio_encode_example(float(label), ids.ctypes.data, vals.ctypes.data, ids.size, buf, cap, int(packed))
    if n < 0:
        raise DataError(_err())
    return buf.raw[:n]


def parse_example(payload: bytes, field_size: int, ids_dtype=np.int64) -> Tuple[float, np.ndarray, np.ndarray]:
    ids = np.empty(field_size, dtype=ids_dtype)
    vals = np.empty(field_size, dtype=np.float32)
    label = C.c_float()
    i32 = int(np.dtype(ids_dtype) == np.int32)
    rc = _lib().dfio_parse_example(payload, len(payload), field_size, ids.ctypes.data, i32, vals.ctypes.data,
                                   C.byref(label))
    if rc != 0:
        raise DataError(_err())
    return label.value, ids, vals


def read_all(path: str, field_size: int, shard_n: int = 1, shard_i: int = 0, ids_dtype=np.int64,
             capacity: Optional[int] = None, verify_crc: bool = True):
    """Parses a whole file into (ids [n,F], vals [n,F], labels [n]) numpy arrays."""
    if capacity is None:
        capacity = (count_records(path, verify_crc=False) + shard_n - 1) // shard_n
    ids = np.empty((capacity, field_size), dtype=ids_dtype)
    vals = np.empty((capacity, field_size), dtype=np.float32)
    labels = np.empty(capacity, dtype=np.float32)
    i32 = int(np.dtype(ids_dtype) == np.int32)
    n = _lib().dfio_read_all(path.encode(), field_size, shard_n, shard_i, capacity, ids.ctypes.data, i32,
                             vals.ctypes.data, labels.ctypes.data, int(verify_crc))
    if n < 0:
        raise DataError(_err())
    return ids[:n], vals[:n], labels[:n]


class TFRecordWriter:
    """Context manager writing framed records (and DeepFM Examples)."""

    def __init__(self, path: str):
        self._lib = _lib()
        self._h = self._lib.dfio_writer_open(path.encode())
        if not self._h:
            raise DataError(_err())

    def write(self, payload: bytes) -> None:
        if self._lib.dfio_writer_write(self._h, payload, len(payload)) != 0:
            raise DataError(_err())

    def write_example(self, label: float, ids, vals) -> None:
        ids = np.ascontiguousarray(ids, dtype=np.int64)
        vals = np.ascontiguousarray(vals, dtype=np.float32)
        if self._lib.dfio_writer_write_example(self._h, float(label), ids.ctypes.data, vals.ctypes.data, ids.size) != 0:
            raise DataError(_err())

    def close(self) -> None:
        if self._h:
            h, self._h = self._h, None
            if self._lib.dfio_writer_close(h) != 0:
                raise DataError(_err())

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def libsvm_to_tfrecord(in_path: str, out_path: str, expect_fields: int = 0) -> int:
    """``<label> <id>:<val> ...`` text -> TFRecord (reference tools/libsvm_to_tfrecord.py:22-61)."""
    n = _lib().dfio_libsvm_to_tfrecord(in_path.encode(), out_path.encode(), expect_fields)
    if n < 0:
        raise DataError(_err())
    return int(n)


def generate_synthetic(out_path: str, n_records: int, field_size: int = 39, feature_size: int = 117581,
                       n_dense: int = 13, seed: int = 1, pos_rate: float = 0.2553, zipf_s: float = 1.05) -> int:
    """Criteo-shaped synthetic TFRecords (statistics modelled on the bundled val.tfrecords)."""
    n = _lib().dfio_generate_synthetic(out_path.encode(), n_records, field_size, feature_size, n_dense, seed,
                                       pos_rate, zipf_s)
    if n < 0:
        raise DataError(_err())
    return int(n)
