"""Fused embedding gather + first-order + FM second-order term on the HOST (csrc/cpu/embed_fm_cpu.cc).

The CPU trainers -- parameter-server workers (the reference's 1-ps-cpu mode, PS.py:196-206) and the eager back-end on a
box without a GPU -- spend more than half of a step in the dozen element-wise PyTorch ops this replaces.  One pass
forward, one pass backward, gradient already de-duplicated per distinct row.  float32 CPU tensors only; anything
else (CUDA tensors, float64 finite-difference checks) takes the plain PyTorch expression in ``models/deepfm.py``,
which is also the oracle this op is tested against.  ``DFM_CPU_NATIVE=0`` turns it off.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from .. import _native

_lib = None


def _get():
    global _lib
    if _lib is None:
        lib = _native.cpu()
        p = C.c_void_p
        lib.dfcpu_embed_fm_fwd.argtypes = [p, p, p, p, C.c_int64, C.c_int64, C.c_int64, p, p, p, p, C.c_int]
        lib.dfcpu_embed_fm_fwd.restype = None
        lib.dfcpu_embed_fm_bwd.argtypes = [p, p, p, p, p, p, p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, p, p, C.c_int]
        lib.dfcpu_embed_fm_bwd.restype = None
        _lib = lib
    return _lib


def usable(w_rows: torch.Tensor, v_rows: torch.Tensor, inverse: torch.Tensor, vals: torch.Tensor) -> bool:
    if os.environ.get("DFM_CPU_NATIVE", "1") == "0":
        return False
    return (not w_rows.is_cuda and not v_rows.is_cuda and not inverse.is_cuda and not vals.is_cuda
            and w_rows.dtype == torch.float32 and v_rows.dtype == torch.float32 and vals.dtype == torch.float32
            and inverse.dtype == torch.int64 and v_rows.dim() == 2 and inverse.dim() == 2 and v_rows.shape[1] <= 512)


class _EmbedFm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, w_rows, v_rows, inverse, vals):
        lib = _get()
        w_rows, v_rows, inverse, vals = (t.contiguous() for t in (w_rows, v_rows, inverse, vals))
        B, F = inverse.shape
        K = v_rows.shape[1]
        y_w = torch.empty(B, dtype=torch.float32)
        y_v = torch.empty(B, dtype=torch.float32)
        e = torch.empty(B, F, K, dtype=torch.float32)
        s = torch.empty(B, K, dtype=torch.float32)
        lib.dfcpu_embed_fm_fwd(w_rows.data_ptr(), v_rows.data_ptr(), inverse.data_ptr(), vals.data_ptr(), B, F, K,
                               y_w.data_ptr(), y_v.data_ptr(), e.data_ptr(), s.data_ptr(), torch.get_num_threads())
        ctx.save_for_backward(inverse, vals, e, s)
        ctx.rows = v_rows.shape[0]
        ctx.mark_non_differentiable(s)
        return y_w, y_v, e, s

    @staticmethod
    def backward(ctx, g_yw, g_yv, g_e, _g_s):
        lib = _get()
        inverse, vals, e, s = ctx.saved_tensors
        B, F = inverse.shape
        K = e.shape[2]
        g_w = torch.zeros(ctx.rows, dtype=torch.float32)
        g_v = torch.zeros(ctx.rows, K, dtype=torch.float32)
        keep = [t.contiguous() if t is not None else None for t in (g_yw, g_yv, g_e)]     # alive across the call
        lib.dfcpu_embed_fm_bwd(inverse.data_ptr(), vals.data_ptr(), e.data_ptr(), s.data_ptr(),
                               *[t.data_ptr() if t is not None else None for t in keep], B, F, K, ctx.rows,
                               g_w.data_ptr(), g_v.data_ptr(), torch.get_num_threads())
        return g_w, g_v, None, None


def embed_fm(w_rows: torch.Tensor, v_rows: torch.Tensor, inverse: torch.Tensor, vals: torch.Tensor):
    """(y_w [B], y_v [B], e [B, F, K]) of the rows ``inverse`` selects; differentiable w.r.t. ``w_rows`` / ``v_rows``."""
    y_w, y_v, e, _ = _EmbedFm.apply(w_rows, v_rows, inverse, vals)
    return y_w, y_v, e
