"""bf16 GEMM on tcgen05/TMEM (csrc/cuda/gemm.cu).

The op that stands in for the cuBLAS SGEMM behind
``tf.contrib.layers.fully_connected`` (reference HVD.py:224-229) when a layer runs
un-fused; the fused MLP kernels specialise the same pipeline.
"""
from __future__ import annotations

import ctypes

import torch

from .. import _native


def _stream_ptr() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, bias: torch.Tensor | None = None, relu: bool = False,
              a_mn_major: bool = False, b_mn_major: bool = False) -> torch.Tensor:
    """C[M,N] fp32 = A @ B^T (+bias)(relu).

    ``a`` is [M,K] (or [K,M] when ``a_mn_major``), ``b`` is [N,K] (or [K,N] when
    ``b_mn_major``); both bf16, contiguous rows, row pitch a multiple of 8 elements.
    """
    assert a.is_cuda and b.is_cuda and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.stride(1) == 1 and b.stride(1) == 1
    if a_mn_major:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn_major:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb, (a.shape, b.shape)
    c = torch.empty((M, N), device=a.device, dtype=torch.float32)
    lib = _native.cuda()
    rc = lib.dfm_gemm_bf16(
        ctypes.c_void_p(a.data_ptr()), ctypes.c_int(a.stride(0)), ctypes.c_int(int(a_mn_major)),
        ctypes.c_void_p(b.data_ptr()), ctypes.c_int(b.stride(0)), ctypes.c_int(int(b_mn_major)),
        ctypes.c_void_p(c.data_ptr()), ctypes.c_int(c.stride(0)),
        ctypes.c_int(M), ctypes.c_int(N), ctypes.c_int(K),
        ctypes.c_void_p(bias.data_ptr() if bias is not None else 0), ctypes.c_int(int(relu)),
        _stream_ptr())
    _native.check(rc, "dfm_gemm_bf16")
    return c
