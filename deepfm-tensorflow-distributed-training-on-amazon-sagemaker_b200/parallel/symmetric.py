"""Symmetric-memory bring-up for the in-kernel collectives.

Each rank's engine allocates one block with an identical layout (dense gradient buffer, averaged
gradient buffer, flag words, sparse inboxes) and exports it as a CUDA IPC handle.  The handles are
exchanged once through ``torch.distributed`` (control plane only -- the role MPI plays for Horovod,
SURVEY X1/X5); after that the kernels in csrc/cuda/comm.cu move every byte themselves with loads and
stores on the mapped peer pointers over NVLink / NVSwitch.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from .. import _native


def attach_engine(engine) -> None:
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() < 2:
        return
    world, rank = dist.get_world_size(), dist.get_rank()
    lib = engine._lib
    lib.dfm_engine_comm_alloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.dfm_engine_comm_attach.argtypes = [C.c_void_p, C.c_void_p]
    handle = C.create_string_buffer(64)
    with torch.cuda.device(engine.device):
        _native.check(lib.dfm_engine_comm_alloc(engine._h, world, rank, handle), "dfm_engine_comm_alloc")
        gathered = [None] * world
        dist.all_gather_object(gathered, bytes(handle.raw))
        blob = b"".join(gathered)
        assert len(blob) == 64 * world
        _native.check(lib.dfm_engine_comm_attach(engine._h, blob), "dfm_engine_comm_attach")
    engine._bufs.clear()          # the gradient buffer moved into the symmetric block
    engine.state_t = engine.buffer("state", dtype="i4")
    dist.barrier()
