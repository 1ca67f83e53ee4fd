"""The fused engine's dropout stream, restated in torch (csrc/cuda/common.cuh: philox4x32, dropout_keep_bits8).

The engine decides "keep" for activation (row b, hidden layer l, column c) at optimizer step t from one Philox4x32-10
block with key = seed and counter = (b, l, c // 8, t): its four 32-bit outputs are eight 16-bit uniforms, and column c
is kept when uniform number (c % 8) is below keep * 65536.  With the same masks the eager fp32 oracle and the fused
bf16 engine differ by rounding only, which is what the precision comparison (scripts/auc_parity.py) wants to isolate.
"""
from __future__ import annotations

import torch

_M0, _M1 = 0xD2511F53, 0xCD9E8D57
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_MASK = 0xFFFFFFFF


_cpu_lib = None


def _native_ok() -> bool:
    """The same stream from csrc/cpu (a tenth of the time of the tensor expression below); DFM_CPU_NATIVE=0: off."""
    global _cpu_lib
    import os
    if os.environ.get("DFM_CPU_NATIVE", "1") == "0":
        return False
    if _cpu_lib is None:
        import ctypes as C
        from .. import _native
        lib = _native.cpu()
        lib.dfcpu_keep_mask.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_int64, C.c_int64, C.c_float, C.c_void_p]
        lib.dfcpu_keep_mask.restype = None
        _cpu_lib = lib
    return True


def _mulhilo(a: int, b: torch.Tensor):
    p = b * a                      # int64 product mod 2^64: the low and the high 32 bits are still the true bits
    return (p >> 32) & _MASK, p & _MASK


def philox4x32(seed: int, c0, c1, c2, c3):
    """Vectorised Philox4x32-10: counters are int64 tensors holding uint32 values; returns four such tensors."""
    k0, k1 = seed & _MASK, (seed >> 32) & _MASK
    for _ in range(10):
        hi0, lo0 = _mulhilo(_M0, c0)
        hi1, lo1 = _mulhilo(_M1, c2)
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & _MASK, lo1, (hi0 ^ c3 ^ k1) & _MASK, lo0
        k0, k1 = (k0 + _W0) & _MASK, (k1 + _W1) & _MASK
    return c0, c1, c2, c3


def keep_mask(seed: int, step: int, layer: int, rows: int, cols: int, keep: float, device=None) -> torch.Tensor:
    """bool [rows, cols]: the engine's dropout keep decisions after hidden layer ``layer`` at 1-based step ``step``."""
    dev = torch.device(device) if device is not None else torch.device("cpu")
    if dev.type == "cpu" and _native_ok():
        out = torch.empty(rows, cols, dtype=torch.uint8)
        _cpu_lib.dfcpu_keep_mask(int(seed) & 0xFFFFFFFFFFFFFFFF, int(step) & _MASK, int(layer) & _MASK, rows, cols,
                                 float(keep), out.data_ptr())
        return out.view(torch.bool)          # bytes 0 / 1: reinterpreted, not converted
    groups = (cols + 7) // 8
    b = torch.arange(rows, device=dev, dtype=torch.int64).view(-1, 1).expand(rows, groups)
    g = torch.arange(groups, device=dev, dtype=torch.int64).view(1, -1).expand(rows, groups)
    o = philox4x32(int(seed), b, torch.full_like(b, int(layer)), g, torch.full_like(b, int(step)))
    thr = int(keep * 65536.0)
    bits = []
    for w in o:                    # word i -> columns 2i (low half) and 2i + 1 (high half)
        bits.append((w & 0xFFFF) < thr)
        bits.append((w >> 16) < thr)
    return torch.stack(bits, dim=-1).reshape(rows, groups * 8)[:, :cols]
