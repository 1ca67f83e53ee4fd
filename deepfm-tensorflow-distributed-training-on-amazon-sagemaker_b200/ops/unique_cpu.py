"""``ids.unique(return_inverse=True)`` for a batch of feature ids on the host (csrc/cpu/unique_cpu.cc).

Same result -- sorted distinct ids and, for every position, its index among them -- without sorting all B x F ids: a
table with one entry per vocabulary row finds the distinct ones (the GPU engine's dedupe kernel works the same way)
and only those are sorted.  Used by the CPU trainers (PS workers, eager back-end on a host); anything else -- CUDA
tensors, vocabularies above 64 M rows (the table would cost more than it saves), ``DFM_CPU_NATIVE=0`` -- takes
``torch.unique``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Tuple

import torch

from .. import _native

_lib = None
_scratch: Dict[int, torch.Tensor] = {}
MAX_TABLE_ROWS = 1 << 26


def unique_inverse(ids: torch.Tensor, feature_size: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """(sorted distinct ids [U] int64, inverse with the shape of ``ids`` int64)."""
    global _lib
    ids = ids.long()
    if (ids.is_cuda or feature_size <= 0 or feature_size > MAX_TABLE_ROWS or ids.numel() == 0
            or os.environ.get("DFM_CPU_NATIVE", "1") == "0"):
        return ids.unique(return_inverse=True)
    if _lib is None:
        lib = _native.cpu()
        lib.dfcpu_unique_inverse.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.dfcpu_unique_inverse.restype = C.c_int64
        _lib = lib
    key = int(feature_size)
    if key not in _scratch:
        _scratch[key] = torch.zeros(key, dtype=torch.int32)
    flat = ids.contiguous().view(-1)
    uniq = torch.empty(flat.numel(), dtype=torch.int64)
    inverse = torch.empty(flat.numel(), dtype=torch.int64)
    u = _lib.dfcpu_unique_inverse(flat.data_ptr(), flat.numel(), key, _scratch[key].data_ptr(), uniq.data_ptr(),
                                  inverse.data_ptr())
    if u < 0:
        bad = int(flat[-1 - u])
        raise IndexError(f"feature id {bad} is outside [0, feature_size = {feature_size})")
    return uniq[:u].clone(), inverse.view(ids.shape)
