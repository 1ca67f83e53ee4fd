"""Batch loader: native producer threads -> ring of pinned host buffers.

The reference's ``input_fn`` (HVD.py:104-161) is
``TFRecordDataset -> shard -> batch(drop_remainder=True) -> parse_example -> repeat -> prefetch``.
``BatchLoader`` is that pipeline as a C++ thread pool (csrc/io/loader.cc) that decodes
straight into page-locked torch tensors, so the trainer can issue one async H2D copy
per tensor with no intermediate staging.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Iterator, List, Optional, Sequence

import torch

from .. import _native
from .tfrecord import DataError, _lib


class _Cfg(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "field_size", "batch_size", "num_epochs", "drop_remainder", "shard_n", "shard_i", "num_workers",
        "num_slots", "ids_i32", "verify_crc", "streaming", "feature_size")]


@dataclass
class HostBatch:
    """Views into one ring slot; valid until the next batch is requested."""
    feat_ids: torch.Tensor   # [n, F] int32|int64, pinned when CUDA is present
    feat_vals: torch.Tensor  # [n, F] float32
    labels: torch.Tensor     # [n]    float32
    count: int
    slab: Optional[torch.Tensor] = None   # the whole slot as bytes: [ids | vals | labels] for batch_size rows
    slot: Optional[int] = None            # ring slot (only needed with BatchLoader.defer_release)


def slab_layout(batch_size: int, field_size: int, ids_dtype: torch.dtype = torch.int32):
    """(total_bytes, offset_vals, offset_labels) of a batch slab -- the layout the fused engine's
    host->device pipeline copies in one piece (dfm_engine_slab_layout)."""
    isz = 4 if ids_dtype == torch.int32 else 8
    bf = batch_size * field_size
    return bf * isz + bf * 4 + batch_size * 4, bf * isz, bf * isz + bf * 4


def alloc_slab(batch_size: int, field_size: int, ids_dtype: torch.dtype = torch.int32, pin_memory: bool = False):
    """One contiguous (optionally pinned) byte buffer plus typed views (ids, vals, labels) into it."""
    total, ov, ol = slab_layout(batch_size, field_size, ids_dtype)
    slab = torch.empty(total, dtype=torch.uint8, pin_memory=pin_memory)
    ids = slab[:ov].view(ids_dtype).view(batch_size, field_size)
    vals = slab[ov:ol].view(torch.float32).view(batch_size, field_size)
    labels = slab[ol:].view(torch.float32)
    return slab, ids, vals, labels


class BatchLoader:
    def __init__(self, paths: Sequence[str], field_size: int, batch_size: int, num_epochs: int = 1,
                 drop_remainder: bool = True, shard_n: int = 1, shard_i: int = 0, num_workers: int = 4,
                 num_slots: int = 4, ids_dtype: torch.dtype = torch.int32, verify_crc: bool = True,
                 streaming: bool = False, pin_memory: Optional[bool] = None, feature_size: int = 0):
        """``feature_size`` > 0: every id must lie in [0, feature_size); a record that violates it ends the stream
        with a DataError naming the file and byte offset (TF's embedding_lookup raises InvalidArgument there)."""
        lib = _lib()
        if not getattr(lib, "_loader_typed", False):
            lib.dfio_loader_create.restype = C.c_void_p
            lib.dfio_loader_create.argtypes = [C.POINTER(_Cfg), C.POINTER(C.c_char_p), C.c_int,
                                               C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
            lib.dfio_loader_next.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
            lib.dfio_loader_release.argtypes = [C.c_void_p, C.c_int]
            lib.dfio_loader_stats.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
            lib.dfio_loader_destroy.argtypes = [C.c_void_p]
            lib._loader_typed = True
        self._lib = lib
        assert ids_dtype in (torch.int32, torch.int64)
        if pin_memory is None:
            pin_memory = torch.cuda.is_available()
        self.field_size, self.batch_size = field_size, batch_size
        self._slabs, self._ids, self._vals, self._labels = [], [], [], []
        for _ in range(num_slots):
            slab, ids, vals, labels = alloc_slab(batch_size, field_size, ids_dtype, pin_memory)
            self._slabs.append(slab); self._ids.append(ids); self._vals.append(vals); self._labels.append(labels)
        cfg = _Cfg(field_size, batch_size, max(1, num_epochs), int(drop_remainder), max(1, shard_n), shard_i,
                   num_workers, num_slots, int(ids_dtype == torch.int32), int(verify_crc), int(streaming),
                   int(feature_size) if 0 < int(feature_size) < 2 ** 31 else 0)
        arr = (C.c_char_p * len(paths))(*[p.encode() for p in paths])
        vp = lambda ts: (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])
        self._h = lib.dfio_loader_create(C.byref(cfg), arr, len(paths), vp(self._ids), vp(self._vals), vp(self._labels))
        if not self._h:
            raise DataError(lib.dfio_last_error().decode())
        self._held: Optional[int] = None
        self.num_slots = num_slots
        # False: a batch is valid until the next one is requested.  True: the consumer returns every slot itself
        # with release(slot) -- the trainer does, once the GPU has finished copying the slab (Trainer.train_batches)
        self.defer_release = False

    def release(self, slot: int) -> None:
        if self._h is not None:
            self._lib.dfio_loader_release(self._h, slot)
        if self._held == slot:
            self._held = None

    def __iter__(self) -> Iterator[HostBatch]:
        return self

    def __next__(self) -> HostBatch:
        if self._h is None:
            raise StopIteration
        if self._held is not None and not self.defer_release:
            self._lib.dfio_loader_release(self._h, self._held)
        self._held = None
        slot, count = C.c_int(), C.c_int()
        r = self._lib.dfio_loader_next(self._h, C.byref(slot), C.byref(count))
        if r == 0:
            raise StopIteration
        if r < 0:
            raise DataError(self._lib.dfio_last_error().decode())
        s, n = slot.value, count.value
        self._held = s
        if n == self.batch_size:     # the common case: the slot's own views, no slicing
            return HostBatch(self._ids[s], self._vals[s], self._labels[s], n, self._slabs[s], s)
        return HostBatch(self._ids[s][:n], self._vals[s][:n], self._labels[s][:n], n, self._slabs[s], s)

    def stats(self):
        r, b, y = C.c_int64(), C.c_int64(), C.c_int64()
        self._lib.dfio_loader_stats(self._h, C.byref(r), C.byref(b), C.byref(y))
        return {"records": r.value, "batches": b.value, "bytes": y.value}

    def close(self) -> None:
        if self._h is not None:
            h, self._h = self._h, None
            self._lib.dfio_loader_destroy(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
