"""Multi-GPU self-check: the fused NVLink collectives against the restated Horovod path on NCCL.

Run by every rank of a ``torch.distributed`` job (``bench.py`` does it before its timed region at N > 1, so the
driver's scaling runs carry the evidence; ``scripts/gpu_dp_check.py`` and ``tests/test_multi_gpu.py`` use it too).
Both trainers start from the same weights and see the same fixed per-rank batches for ``steps`` optimizer steps of
plain gradient descent (no slot state, dropout off, so the comparison is sharp):

* ``replicas_identical``            -- after the fused steps every parameter tensor is bit-identical on all ranks
  (each sum is formed by exactly one rank and published, csrc/cuda/comm.cu);
* ``worst_rel_update_err_vs_nccl``  -- max over parameter tensors of ||du_fused - du_eager|| / ||du_eager||, where du is
  the total update and the eager arm is fp32 PyTorch autograd + ``dist.all_reduce`` / ``all_gather``
  (hvd.DistributedOptimizer semantics, HVD.py:296).  The fused arm computes the MLP in bf16, hence a few percent.
"""
from __future__ import annotations

from typing import Dict, Sequence

import torch
import torch.distributed as dist


def fixed_batches(rank: int, steps: int, B: int, F: int, V: int, seed: int = 100):
    g = torch.Generator().manual_seed(seed + rank)
    out = []
    for _ in range(steps):
        ids = torch.randint(1, V, (B, F), generator=g, dtype=torch.int32)
        ids[:, :13] = torch.arange(1, 14, dtype=torch.int32)       # Criteo numeric fields: same ids on every rank
        vals = torch.rand(B, F, generator=g)
        vals[:, 13:] = 1.0
        labels = (torch.rand(B, generator=g) < 0.3).float()
        out.append((ids, vals, labels))
    return out


def fused_vs_nccl(device: torch.device, field_size: int = 39, feature_size: int = 117581, embedding_size: int = 32,
                  deep_layers: Sequence[int] = (256, 128, 64), batch: int = 2048, steps: int = 4, lr: float = 0.05,
                  optimizer: str = "GD", l2_reg: float = 1e-3, cuda_graph: bool = True) -> Dict[str, object]:
    from ..data.loader import HostBatch, alloc_slab
    from ..models.deepfm import DeepFM, DeepFMConfig
    from ..runtime.trainer import Trainer, TrainerOptions

    rank, world = dist.get_rank(), dist.get_world_size()
    cfg = DeepFMConfig(field_size, feature_size, embedding_size, tuple(deep_layers), (1.0,) * len(deep_layers),
                       l2_reg=l2_reg)
    init = DeepFM(cfg, seed=3)
    with torch.no_grad():
        init.fm_v.mul_(20)
        init.fm_w.mul_(20)
    data = fixed_batches(rank, steps, batch, field_size, feature_size)
    w0 = {k: v.detach().clone().to(device) for k, v in init.tf_variables().items()}

    fused = Trainer(cfg, TrainerOptions(batch_size=batch, optimizer=optimizer, learning_rate=lr, backend="fused", seed=3,
                                        cuda_graph=cuda_graph), device=device, init_model=init)
    for ids, vals, labels in data:
        slab, sid, sv, sl = alloc_slab(batch, field_size, torch.int32, pin_memory=True)
        sid.copy_(ids); sv.copy_(vals); sl.copy_(labels)
        fused.train_step_host(slab)
        fused.last_loss(wait_current=True)
    torch.cuda.synchronize(device)
    fv = {k: v.detach().clone() for k, v in fused.variables().items()}
    info = {"provider": getattr(fused.comm_provider, "name", None),
            "multicast": int(fused.engine._lib.dfm_engine_comm_multicast(fused.engine._h)) if world > 1 else 0}
    fused.close()

    eager = Trainer(cfg, TrainerOptions(batch_size=batch, optimizer=optimizer, learning_rate=lr, backend="eager", seed=3),
                    device=device, init_model=DeepFM(cfg, seed=3))
    with torch.no_grad():
        for k, dst in eager.model.tf_variables().items():
            dst.copy_(w0[k])
    for ids, vals, labels in data:
        eager.train_batch(HostBatch(ids, vals, labels, batch))
    ev = {k: v.detach() for k, v in eager.variables().items()}

    identical = True
    for k, v in fv.items():
        ref = v.clone()
        dist.broadcast(ref, src=0)
        identical = identical and bool(torch.equal(ref, v))
    flag = torch.tensor([1 if identical else 0], device=device, dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    errs = {}
    for k, v in fv.items():
        du_f, du_e = v - w0[k], ev[k] - w0[k]
        errs[k] = float((du_f - du_e).norm() / (du_e.norm() + 1e-12))
    worst = torch.tensor([max(errs.values())], device=device, dtype=torch.float64)
    dist.all_reduce(worst, op=dist.ReduceOp.MAX)
    return {"replicas_identical": bool(flag.item()), "worst_rel_update_err_vs_nccl": float(worst.item()),
            "rel_update_err": {k: round(e, 4) for k, e in errs.items()}, "steps": steps, "optimizer": optimizer, **info}
