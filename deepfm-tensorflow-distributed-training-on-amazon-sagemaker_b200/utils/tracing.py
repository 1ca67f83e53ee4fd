"""NVTX ranges for Nsight Systems / Compute timelines (the reference only has NCCL_DEBUG and Estimator's
steps/sec log; its launchers switch SageMaker's profiler off, NB-HVD:130-131).

``DFM_NVTX=1`` turns the ranges on: ``train_step``, ``evaluate``, ``predict``, ``checkpoint``, ``export`` and the loader's
``next_batch``.  Off (the default) every call is a no-op, so the 58 us training step pays nothing for them.
The kernels carry their own evidence independently of this: ``%globaltimer`` stamps in ``StepState::kstamp`` and the
collective kernels' phase stamps (``DFM_BENCH_TIMELINE=1`` / ``DFM_BENCH_PHASES=1 python bench.py``).
"""
from __future__ import annotations

import contextlib
import os

ENABLED = os.environ.get("DFM_NVTX", "0") not in ("", "0")


def push(name: str) -> None:
    if ENABLED:
        import torch
        torch.cuda.nvtx.range_push(name)


def pop() -> None:
    if ENABLED:
        import torch
        torch.cuda.nvtx.range_pop()


@contextlib.contextmanager
def nvtx_range(name: str):
    if not ENABLED:
        yield
        return
    import torch
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()
