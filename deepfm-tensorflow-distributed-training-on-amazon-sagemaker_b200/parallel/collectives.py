"""Stand-alone collectives on symmetric memory, and the collective sweep (BASELINE.json config 5).

``SymmetricAllReduce`` exposes the dense half of the training step's collective kernel on its own: a sum all-reduce of a
buffer that lives in symmetric memory, done by one kernel per rank that reduces its chunk over all replicas
(``multimem.ld_reduce`` inside the NVSwitch where available, peer loads otherwise) and publishes it (``multimem.st`` /
peer stores) -- no NCCL call (csrc/cuda/comm.cu, ``allreduce_kernel``).  It stands where Horovod's fusion-buffer
``ncclAllReduce`` does (HVD.py:296).

``sweep`` times it against ``torch.distributed.all_reduce`` (NCCL) over 4 KB .. 256 MB, and the fused sparse exchange
(records to owners, merge, lazy Adam, rows into every replica) against the restated Horovod path
(all_gather of ids and rows + merge + row update) over 1 K .. 1 M distinct ids per rank.  Device-timed, max over ranks.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List

import torch
import torch.distributed as dist

from .. import _native
from .symmetric import SymmetricProvider, make_provider

_FLAG_WORDS = 2 * 8 * 160


class SymmetricAllReduce:
    def __init__(self, max_elems: int, device: torch.device, provider: SymmetricProvider = None):
        self.provider = provider or make_provider(device)
        self.world, self.rank = self.provider.world, self.provider.rank
        self.device = device
        n = (max_elems + 3) // 4 * 4
        self.data = self.provider.alloc((n,), torch.float32, device)
        self.flags = self.provider.alloc((_FLAG_WORDS,), torch.int32, device)
        dptrs, self._mc = self.provider.exchange(self.data)
        fptrs, _ = self.provider.exchange(self.flags)
        self._dptrs = (C.c_void_p * self.world)(*[C.c_void_p(p) for p in dptrs])
        self._fptrs = (C.c_void_p * self.world)(*[C.c_void_p(p) for p in fptrs])
        self._epoch = 0
        lib = _native.cuda()
        lib.dfm_allreduce_f32.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_longlong, C.c_uint,
                                          C.c_void_p]
        self._lib = lib
        dist.barrier()

    @property
    def uses_multicast(self) -> bool:
        return bool(self._mc)

    def all_reduce_(self, n_elems: int) -> torch.Tensor:
        """Sums ``self.data[:n_elems]`` over all ranks, in place in every rank's copy.  Collective."""
        n = (n_elems + 3) // 4 * 4
        self._epoch += 1
        _native.check(self._lib.dfm_allreduce_f32(self.world, self.rank, self._dptrs, C.c_void_p(self._mc) if self._mc else None,
                                                  self._fptrs, n, self._epoch,
                                                  C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)),
                      "dfm_allreduce_f32")
        return self.data[:n_elems]


def _time_us(fn, iters: int, device, pre=None) -> float:
    """Mean device time of fn() in microseconds (CUDA events, max over ranks); pre() runs untimed before each call."""
    for _ in range(3):
        if pre:
            pre()
        fn()
    torch.cuda.synchronize(device)
    total = 0.0
    if pre is None:
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize(device)
        total = e0.elapsed_time(e1) * 1e3 / iters
    else:
        for _ in range(iters):
            pre()
            torch.cuda.synchronize(device)
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize(device)
            total += e0.elapsed_time(e1) * 1e3 / iters
    t = torch.tensor([total], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sweep_dense(device: torch.device, sizes_bytes: List[int]) -> List[Dict[str, float]]:
    world = dist.get_world_size()
    ar = SymmetricAllReduce(max(sizes_bytes) // 4, device)
    rows = []
    for nbytes in sizes_bytes:
        n = nbytes // 4
        ar.data[:n].fill_(1.0)
        torch.cuda.synchronize(device)
        out = ar.all_reduce_(n)
        torch.cuda.synchronize(device)
        assert float(out[0]) == float(world) and float(out[n - 1]) == float(world), "fused all-reduce: wrong sum"
        iters = 200 if nbytes <= (1 << 22) else 20
        fused = _time_us(lambda: ar.all_reduce_(n), iters, device)
        x = torch.ones(n, device=device)
        nccl = _time_us(lambda: dist.all_reduce(x), iters, device)
        # bus bandwidth as NCCL reports it: 2 (N-1)/N * bytes / time
        bw = lambda us: 2.0 * (world - 1) / world * nbytes / (us * 1e-6) / 1e9
        rows.append({"bytes": nbytes, "fused_us": round(fused, 2), "nccl_us": round(nccl, 2),
                     "fused_busbw_gbs": round(bw(fused), 1), "nccl_busbw_gbs": round(bw(nccl), 1)})
    return rows


def sweep_sparse(device: torch.device, counts: List[int], feature_size: int = 4_000_000, Ke: int = 32) -> List[Dict[str, float]]:
    from ..models.deepfm import DeepFMConfig
    from ..runtime.engine import FusedEngine
    from ..utils.optimizers import TFOptimizer
    from .symmetric import attach_engine
    from .sync_dp import sparse_allgather_mean
    world, rank = dist.get_world_size(), dist.get_rank()
    nmax = max(counts)
    cfg = DeepFMConfig(field_size=1, feature_size=feature_size, embedding_size=Ke, deep_layers=(16,), dropout_keep=(1.0,),
                       l2_reg=1e-4)
    prov = make_provider(device)
    eng = FusedEngine(cfg, nmax, optimizer="Adam", lr=1e-3, device=device, provider=prov)
    attach_engine(eng, prov)
    lib = eng._lib
    lib.dfm_engine_exchange_prepare.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.dfm_engine_exchange_run.argtypes = [C.c_void_p, C.c_void_p]
    stream = lambda: C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
    # the restated path keeps its own table and slots
    tab_w = torch.zeros(feature_size, device=device)
    tab_v = torch.zeros(feature_size, Ke, device=device)
    opt = TFOptimizer("Adam", 1e-3)
    rows = []
    g = torch.Generator(device=device).manual_seed(1234 + rank)
    for n in counts:
        ids = torch.randperm(feature_size, device=device, generator=g)[:n].to(torch.int32).contiguous()   # distinct per rank
        ids64 = ids.long()
        gv = torch.ones(n, Ke, device=device)
        gw = torch.ones(n, device=device)
        pre = lambda: _native.check(lib.dfm_engine_exchange_prepare(eng._h, ids.data_ptr(), n, stream()), "prepare")
        run = lambda: _native.check(lib.dfm_engine_exchange_run(eng._h, stream()), "run")

        def nccl_path():
            u, mw, mv = sparse_allgather_mean(ids64, gw, gv, feature_size)
            opt.apply_rows("fm_w", tab_w, u, mw + 1e-4 * tab_w.index_select(0, u))
            opt.apply_rows("fm_v", tab_v, u, mv + 1e-4 * tab_v.index_select(0, u))
        iters = 30 if n <= 100_000 else 8
        fused = _time_us(run, iters, device, pre=pre)
        nccl = _time_us(nccl_path, iters, device)
        rows.append({"unique_ids_per_rank": n, "fused_us": round(fused, 2), "nccl_us": round(nccl, 2),
                     "speedup": round(nccl / fused, 2)})
    eng.close()
    return rows
