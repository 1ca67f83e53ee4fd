"""Distinct feature ids of a batch with compact slots (csrc/cuda/unique.cu).

The sparse-gradient path reduces the B*F per-position row gradients of a batch into one row per
distinct id before anything is exchanged or applied; ``unique_slots`` is that bookkeeping step as a
standalone op (the reference leaves duplicates in its IndexedSlices, HVD.py:197,201).
"""
from __future__ import annotations

import ctypes as C
from typing import Tuple

import torch

from .. import _native


def _lib():
    lib = _native.cuda()
    if not getattr(lib, "_unique_typed", False):
        lib.dfm_unique_slots_workspace_bytes.restype = C.c_longlong
        lib.dfm_unique_slots_workspace_bytes.argtypes = [C.c_longlong]
        lib.dfm_unique_slots_init.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p]
        lib.dfm_unique_slots.argtypes = [C.c_void_p, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p]
        lib._unique_typed = True
    return lib


class UniqueSlots:
    """Reusable workspace for ids in ``[0, feature_size)`` on one device."""

    def __init__(self, feature_size: int, device=None):
        self.V = int(feature_size)
        self.device = torch.device(device or "cuda")
        lib = _lib()
        self.workspace = torch.empty(lib.dfm_unique_slots_workspace_bytes(self.V), dtype=torch.uint8, device=self.device)
        _native.check(lib.dfm_unique_slots_init(self.workspace.data_ptr(), self.V, self._stream()), "dfm_unique_slots_init")

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def __call__(self, ids: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """ids int32 (any shape) -> (uniq [n] with the first ``count`` valid, pos_slot like ids, count [1])."""
        assert ids.is_cuda and ids.dtype == torch.int32 and ids.is_contiguous()
        n = ids.numel()
        uniq = torch.empty(n, dtype=torch.int32, device=ids.device)
        pos = torch.empty_like(ids)
        count = torch.empty(1, dtype=torch.int32, device=ids.device)
        _native.check(_lib().dfm_unique_slots(ids.data_ptr(), n, self.V, self.workspace.data_ptr(), uniq.data_ptr(),
                                              pos.data_ptr(), count.data_ptr(), self._stream()), "dfm_unique_slots")
        return uniq, pos, count
