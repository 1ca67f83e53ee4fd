"""Synchronous data parallelism (the reference's Horovod mode, HVD.py:296).

``hvd.DistributedOptimizer`` averages every gradient across ranks each step: dense tensors
through an all-reduce, ``IndexedSlices`` (embedding gradients) through an all-gather of
values and indices followed by a division by the world size.

Two implementations live here:

* the *restated* path on ``torch.distributed`` collectives (NCCL on GPUs, gloo on CPUs):
  :func:`allreduce_mean_` and :func:`sparse_allgather_mean`.  It backs the eager back-end,
  the CPU tests, and is the measured baseline ("a path that only calls NCCL is the baseline");
* the *fused* path: :func:`attach_fused_collectives` maps symmetric buffers into every rank and
  switches the engine to kernels that read/write peer memory over NVLink directly
  (csrc/cuda/comm.cu), with no NCCL call inside the step.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist


def allreduce_mean_(grads: Sequence[torch.Tensor]) -> None:
    """In-place average of dense gradients through ONE flattened all-reduce (Horovod's fusion buffer)."""
    world = dist.get_world_size()
    if world == 1 or not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    flat.div_(world)
    o = 0
    for g in grads:
        n = g.numel()
        g.copy_(flat[o:o + n].view_as(g))
        o += n


def sparse_allgather_mean(ids: torch.Tensor, gw: torch.Tensor, gv: torch.Tensor,
                          feature_size: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """All-gather of (distinct ids, first-order grads [u], factor grads [u, K]) from every rank, merged
    by id (sum) and divided by the world size.  Returns (ids, gw, gv) with distinct ids, identical on
    every rank and in the same order, so replicas apply bit-identical updates."""
    world = dist.get_world_size()
    if world == 1:
        return ids, gw, gv
    dev = ids.device
    n = torch.tensor([ids.numel()], device=dev, dtype=torch.int64)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n)
    counts = [int(c.item()) for c in counts]
    cap = max(counts)
    K = gv.shape[1]
    pad_ids = torch.zeros(cap, device=dev, dtype=torch.int64)
    pad_ids[:ids.numel()] = ids
    pad_g = torch.zeros(cap, K + 1, device=dev, dtype=gv.dtype)
    pad_g[:ids.numel(), :K] = gv
    pad_g[:ids.numel(), K] = gw
    all_ids = [torch.empty_like(pad_ids) for _ in range(world)]
    all_g = [torch.empty_like(pad_g) for _ in range(world)]
    dist.all_gather(all_ids, pad_ids)
    dist.all_gather(all_g, pad_g)
    cat_ids = torch.cat([a[:c] for a, c in zip(all_ids, counts)])
    cat_g = torch.cat([a[:c] for a, c in zip(all_g, counts)])
    uniq, inv = cat_ids.unique(return_inverse=True)          # sorted -> same order on every rank
    merged = torch.zeros(uniq.numel(), K + 1, device=dev, dtype=cat_g.dtype)
    merged.index_add_(0, inv, cat_g)                          # rank-major order: deterministic
    merged.div_(world)
    return uniq, merged[:, K].contiguous(), merged[:, :K].contiguous()


def attach_fused_collectives(trainer) -> None:
    """Publishes this rank's symmetric block and embedding tables to every rank and enables the in-kernel NVLink
    collectives of the fused engine."""
    from .symmetric import attach_engine
    attach_engine(trainer.engine, trainer.comm_provider)


def gather_owner_sharded_slots(engine) -> None:
    """Multi-GPU lazy update keeps the optimizer slots of table row ``id`` current on rank ``id % world`` only.
    Before a checkpoint (or anything else that reads all slots on one rank) every rank fetches the rows it does
    not own from their owners: one broadcast of a strided slice per owner and slot tensor."""
    world = dist.get_world_size()
    if world == 1 or engine.table_update_mode != "lazy":
        return
    for t in (engine.fm_w_s0, engine.fm_w_s1, engine.fm_v_s0, engine.fm_v_s1):
        for owner in range(world):
            piece = t[owner::world].contiguous()
            dist.broadcast(piece, src=owner)
            t[owner::world] = piece


def chief_decides(flag: bool, device) -> bool:
    """COLLECTIVE: rank 0's ``flag``, on every rank.

    For any decision that changes how many steps a rank launches -- stop after a time limit, checkpoint now -- taken
    from a local clock or file system: every step of the fused path contains collective kernels that wait for the same
    step on every peer, so ranks that decided differently hang until the communication watchdog fires.  Call it at the
    same step count on all ranks (``train.save_is_due`` and ``bench.measure_pipeline`` do, every N steps); it drains
    the launching stream, so not every step."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() < 2:
        return bool(flag)
    t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=device)
    dist.broadcast(t, src=0)
    return bool(t.item())
