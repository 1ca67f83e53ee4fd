"""Symmetric-memory bring-up for the in-kernel collectives (csrc/cuda/comm.cu).

The kernels need three kinds of buffer to be addressable from every rank: one *symmetric block* per rank (dense
gradient vector, flag words, sparse inboxes), and every rank's replica of the two embedding tables ``fm_v`` /
``fm_w`` (the owner of a row stores its updated value straight into every replica).  A *provider* allocates such
buffers and exchanges their addresses once, through ``torch.distributed`` -- the control-plane role MPI plays for
Horovod.  After that every byte moves through loads, stores and ``multimem`` instructions the
kernels issue themselves over NVLink / NVSwitch.

Providers:

* :class:`TorchSymmProvider` -- ``torch.distributed._symmetric_memory`` (cuMem allocations whose handles are passed
  between the processes; where the fabric supports NVLS it also binds them to a multicast object, which gives the
  kernels one address that stores to / reduces over all replicas);
* :class:`IpcProvider` -- plain ``cudaMalloc`` + CUDA IPC handles (no multicast; the kernels loop over the peers);
* :class:`LocalProvider` -- N virtual ranks inside one process on one GPU (tests / profiling, see
  ``runtime.engine.EmulatedRanks``).

``DFM_COMM_PROVIDER=symm|ipc`` forces one; the default tries symmetric memory first and falls back to IPC if any
rank cannot set it up.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from .. import _native


class _DevMem:
    def __init__(self, ptr: int, n: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


class SymmetricProvider:
    name = "none"
    world = 1
    rank = 0

    def alloc(self, shape: Sequence[int], dtype: torch.dtype, device: torch.device) -> torch.Tensor:
        """A zero-filled tensor that :meth:`exchange` can publish to the other ranks."""
        raise NotImplementedError

    def exchange(self, t: torch.Tensor) -> Tuple[List[int], int]:
        """Collective: (device pointer of this tensor on every rank -- peers' through a mapping --, multicast
        address of all of them or 0)."""
        raise NotImplementedError


class TorchSymmProvider(SymmetricProvider):
    name = "torch_symm_mem"

    def __init__(self, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self._sm = symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self._handles = []          # keep the rendezvous handles (and with them the peer mappings) alive

    def alloc(self, shape, dtype, device):
        t = self._sm.empty(*shape, dtype=dtype, device=device)
        t.zero_()
        return t

    def exchange(self, t):
        hdl = self._sm.rendezvous(t, self.group)
        self._handles.append(hdl)
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        off = t.data_ptr() - ptrs[self.rank]          # position of the tensor inside the rendezvoused allocation
        if off < 0 or off + t.numel() * t.element_size() > int(hdl.buffer_size):
            raise RuntimeError("symmetric memory: tensor lies outside its rendezvoused buffer")
        mc = int(hdl.multicast_ptr) if getattr(hdl, "has_multicast_support", False) and hdl.multicast_ptr else 0
        return [p + off for p in ptrs], (mc + off if mc else 0)


class IpcProvider(SymmetricProvider):
    name = "cuda_ipc"

    def __init__(self, group=None):
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        lib = _native.cuda()
        lib.dfm_dev_alloc.argtypes = [C.c_longlong, C.POINTER(C.c_void_p)]
        lib.dfm_dev_free.argtypes = [C.c_void_p]
        lib.dfm_ipc_export.argtypes = [C.c_void_p, C.c_void_p]
        lib.dfm_ipc_open.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        self._lib = lib

    def alloc(self, shape, dtype, device):
        n = 1
        for d in shape:
            n *= int(d)
        esize = torch.empty(0, dtype=dtype).element_size()
        p = C.c_void_p()
        with torch.cuda.device(device):
            # a cudaMalloc of its own: IPC handles name whole allocations (never freed: peers may hold mappings)
            _native.check(self._lib.dfm_dev_alloc(max(n * esize, 256), C.byref(p)), "dfm_dev_alloc")
        flat = torch.as_tensor(_DevMem(p.value, max(n * esize, 256), "|u1"), device=device)
        return flat[:n * esize].view(dtype).view(*shape)

    def exchange(self, t):
        handle = C.create_string_buffer(64)
        _native.check(self._lib.dfm_ipc_export(C.c_void_p(t.data_ptr()), handle), "dfm_ipc_export")
        gathered: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(gathered, bytes(handle.raw), group=self.group)
        ptrs = []
        for r, h in enumerate(gathered):
            if r == self.rank:
                ptrs.append(t.data_ptr())
                continue
            p = C.c_void_p()
            _native.check(self._lib.dfm_ipc_open(C.create_string_buffer(h, 64), C.byref(p)), "dfm_ipc_open")
            ptrs.append(p.value)
        return ptrs, 0


class LocalProvider(SymmetricProvider):
    """Single-process stand-in: buffers are ordinary tensors (all 'ranks' live on one GPU)."""
    name = "local"

    def alloc(self, shape, dtype, device):
        return torch.zeros(*shape, dtype=dtype, device=device)


def _agree(ok: bool, device: torch.device) -> bool:
    t = torch.tensor([1 if ok else 0], device=device, dtype=torch.int32)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


def make_provider(device: torch.device) -> Optional[SymmetricProvider]:
    """The provider every rank agrees on, or None when there is a single rank."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() < 2:
        return None
    want = os.environ.get("DFM_COMM_PROVIDER", "auto")
    if want in ("auto", "symm"):
        prov, ok, why = None, True, ""
        try:
            with torch.cuda.device(device):
                prov = TorchSymmProvider()
                probe = prov.alloc((1024,), torch.float32, device)
                ptrs, _ = prov.exchange(probe)
                ok = len(ptrs) == prov.world and all(ptrs)
        except Exception as e:       # noqa: BLE001 -- any failure means "use the other provider"
            ok, why = False, f"{type(e).__name__}: {e}"
        if _agree(ok, device):
            return prov
        if want == "symm":
            raise RuntimeError(f"DFM_COMM_PROVIDER=symm: torch symmetric memory is not usable here ({why})")
        if dist.get_rank() == 0:
            print(f"[symmetric] torch symmetric memory unavailable ({why or 'on another rank'}); using CUDA IPC")
    return IpcProvider()


def attach_engine(engine, provider: SymmetricProvider) -> None:
    """Allocates the engine's symmetric block, publishes it and the tables, and switches the engine to the
    in-kernel collectives."""
    lib = engine._lib
    lib.dfm_engine_comm_bytes.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_longlong)]
    lib.dfm_engine_comm_attach.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 8
    world, rank = provider.world, provider.rank
    nbytes = C.c_longlong()
    _native.check(lib.dfm_engine_comm_bytes(engine._h, world, C.byref(nbytes)), "dfm_engine_comm_bytes")
    with torch.cuda.device(engine.device):
        block = provider.alloc((nbytes.value,), torch.uint8, engine.device)
        torch.cuda.synchronize(engine.device)
        tables = [block, engine.fm_v, engine.fm_w]
        if engine.table_update_mode == "dense_exact":
            engine.grad_table = provider.alloc((engine.cfg.feature_size, engine.cfg.embedding_size + 8), torch.float32,
                                               engine.device)
            tables.append(engine.grad_table)
        views = [provider.exchange(t) for t in tables]
        if len(views) == 3:
            views.append(([0] * world, 0))
        arrs = [(C.c_void_p * world)(*[C.c_void_p(p) for p in ptrs]) for ptrs, _ in views]
        mcs = [C.c_void_p(mc) if mc else None for _, mc in views]
        has_grad = engine.table_update_mode == "dense_exact"
        _native.check(lib.dfm_engine_comm_attach(engine._h, world, rank, arrs[0], mcs[0], arrs[1], mcs[1], arrs[2], mcs[2],
                                                 arrs[3] if has_grad else None, mcs[3] if has_grad else None),
                      "dfm_engine_comm_attach")
    engine._comm_block = block          # keep alive
    engine._comm_provider = provider
    engine._bufs.clear()                # the sparse-gradient buffers were re-allocated with one segment per owner
    engine.state_t = engine.buffer("state", dtype="i4")
    engine.world, engine.rank = world, rank
    dist.barrier()
