"""Trainer: the Estimator-shaped public API (train step / evaluate / predict / checkpoint / export).

It is what the reference gets from ``tf.estimator.Estimator(model_fn, model_dir, params)``
(HVD.py:402-415) plus ``hvd.DistributedOptimizer`` (HVD.py:296) and the broadcast hook
(HVD.py:418), with two interchangeable compute back-ends:

* ``fused``  -- :class:`FusedEngine`: the sm_100a kernels, CUDA-graph replay, double-buffered
  pinned-host input pipeline, in-kernel NVLink collectives when world_size > 1;
* ``eager``  -- plain PyTorch autograd over :class:`DeepFM` with the TF-formula optimizers and
  ``torch.distributed`` collectives (dense all_reduce + sparse all_gather: the restated
  Horovod path).  It is the CPU implementation, the oracle, and the measured baseline.

Semantics shared by both: the L2 penalty on fm_w / fm_v is never communicated (it is identical on
every rank: mean_r(g_r + l2 V) = mean_r(g_r) + l2 V) and is applied inside the table update, either
to the touched rows (``lazy``) or to every row (``dense_exact``, the reference's exact math).
"""
from __future__ import annotations

import os
import time
from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, Iterable, Iterator, List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

from ..data.loader import HostBatch, slab_layout
from ..models.deepfm import DeepFM, DeepFMConfig
from ..utils import tracing
from ..utils.metrics import StreamingAUC
from ..utils.optimizers import TFOptimizer, canonical_optimizer


@dataclass
class TrainerOptions:
    batch_size: int = 64
    optimizer: str = "Adam"
    learning_rate: float = 5e-4
    table_update_mode: str = "lazy"       # lazy | dense_exact
    backend: str = "auto"                 # auto | fused | eager
    seed: int = 0
    cuda_graph: bool = True
    scale_lr_by_world: bool = True        # HVD.py:171 (sync-DP); the PS script does not scale (PS.py:180)
    log_steps: int = 0
    dropout_stream: str = "auto"          # eager back-end: "torch" (its own RNG) | "engine" (the fused engine's Philox
                                          # masks for the same seed and step: a rounding-only comparison, utils/philox.py)
                                          # | "auto": engine on the host, where the native mask generator costs a tenth
                                          # of torch's scalar dropout; torch on a GPU


def _world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class Trainer:
    def __init__(self, cfg: DeepFMConfig, opts: TrainerOptions, device: Optional[torch.device] = None,
                 init_model: Optional[DeepFM] = None):
        self.cfg, self.opts = cfg, opts
        self.rank, self.world = _world()
        self.device = torch.device(device) if device is not None else torch.device(
            "cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        self.lr = opts.learning_rate * (self.world if opts.scale_lr_by_world else 1)
        self.global_step = 0
        self.backend = self._pick_backend()
        # identical initial weights on every rank: same seed, then an explicit rank-0 broadcast (HVD.py:418)
        model = init_model if init_model is not None else DeepFM(cfg, seed=opts.seed)
        self.uses_graph = False
        self._loss_ring = None
        if self.backend == "fused":
            from .engine import FusedEngine
            self.comm_provider = None
            if self.world > 1:
                from ..parallel.symmetric import make_provider
                self.comm_provider = make_provider(self.device)
            self.engine = FusedEngine(cfg, opts.batch_size, optimizer=opts.optimizer, lr=self.lr,
                                      table_update_mode=opts.table_update_mode, seed=opts.seed, device=self.device,
                                      provider=self.comm_provider)
            self.engine.load_from_module(model)
            self.model = None
            self._host_model = model          # kept as a CPU template for export / variable shapes
            if self.world > 1:
                from ..parallel.sync_dp import attach_fused_collectives
                attach_fused_collectives(self)
            self.broadcast_parameters()
            if opts.cuda_graph:
                self.engine.capture(opts.batch_size)
                self.uses_graph = True
                self._eng_loss = self.engine.loss_ring()     # pinned, written by a node of the step graph
            pin = True
            self._loss_ring = [torch.zeros(1, dtype=torch.float32, pin_memory=pin) for _ in range(2)]
            self._loss_ev = [torch.cuda.Event(), torch.cuda.Event()]
            self._loss_src = [None, None]
            self._loss_cnt = [opts.batch_size, opts.batch_size]
            self._loss_idx = 0
            self._issued = 0
        else:
            self.engine = None
            self.model = model.to(self.device)
            self.optimizer = TFOptimizer(opts.optimizer, self.lr)
            self.broadcast_parameters()
            self._last_loss_t = None
            self._prev_loss_t = None

    # ------------------------------------------------------------------ set-up
    def _pick_backend(self) -> str:
        want = self.opts.backend
        if want == "eager":
            return "eager"
        from .engine import fused_engine_supported
        why = None if self.device.type == "cuda" else "no CUDA device"
        why = why or fused_engine_supported(self.cfg)
        if why is None and torch.cuda.get_device_capability(self.device)[0] < 10:
            why = "the kernels are built for sm_100a only"
        if why is None:
            return "fused"
        if want == "fused":
            raise RuntimeError(f"--backend fused is not possible here: {why}")
        if self.rank == 0:
            print(f"[trainer] using the eager PyTorch back-end: {why}")
        return "eager"

    def broadcast_parameters(self, root: int = 0) -> None:
        """hvd.BroadcastGlobalVariablesHook(0) (HVD.py:418): weights, optimizer slots, global step."""
        if self.world <= 1:
            return
        tensors = list(self.state_tensors().values())
        for t in tensors:
            dist.broadcast(t, src=root)
        step = torch.tensor([self.global_step], device=self.device, dtype=torch.int64)
        dist.broadcast(step, src=root)
        self.global_step = int(step.item())
        if self.backend == "fused":
            self.engine.sync_weights()
            self.engine.set_state(self.opts.batch_size, train=True, step=self.global_step)

    # ------------------------------------------------------------------ state
    def variables(self) -> "OrderedDict[str, torch.Tensor]":
        return self.engine.variables() if self.backend == "fused" else self.model.tf_variables()

    def state_tensors(self) -> "OrderedDict[str, torch.Tensor]":
        """Everything a checkpoint holds except global_step, under TF names."""
        d: "OrderedDict[str, torch.Tensor]" = OrderedDict(self.variables())
        if self.backend == "fused":
            d.update(self.engine.slot_variables())
        else:
            for name, p in self.model.trainable_tf_variables().items():
                self.optimizer.ensure_slots(name, p.data)
            d.update(self.optimizer.slots)
        return d

    def prepare_checkpoint(self) -> None:
        """COLLECTIVE (every rank calls it, before the chief calls :meth:`checkpoint_tensors`): drains the step
        queue and, on the fused multi-GPU path, brings the optimizer slots of the table rows other ranks own to this
        rank (the lazy update keeps them current on the owner only, csrc/cuda/comm.cu)."""
        if self.backend == "fused":
            torch.cuda.synchronize(self.device)
            if self.world > 1:
                from ..parallel.sync_dp import gather_owner_sharded_slots
                gather_owner_sharded_slots(self.engine)
                torch.cuda.synchronize(self.device)

    def checkpoint_tensors(self) -> Dict[str, np.ndarray]:
        if self.backend == "fused":
            torch.cuda.synchronize(self.device)
        with tracing.nvtx_range("checkpoint"):
            out = {k: v.detach().cpu().numpy().copy() for k, v in self.state_tensors().items()}
        if canonical_optimizer(self.opts.optimizer) == "adam":
            # tf.train.AdamOptimizer's non-slot variables: beta^t of the NEXT update (betas are the TF defaults the
            # engine and the eager optimizer are built with)
            b1, b2 = self.adam_betas()
            out["beta1_power"] = np.asarray(b1 ** (self.global_step + 1), dtype=np.float32)
            out["beta2_power"] = np.asarray(b2 ** (self.global_step + 1), dtype=np.float32)
        return out

    def adam_betas(self) -> Tuple[float, float]:
        if self.backend == "fused":
            return float(self.engine.beta1), float(self.engine.beta2)
        return float(getattr(self.optimizer, "beta1", 0.9)), float(getattr(self.optimizer, "beta2", 0.999))

    def load_checkpoint_tensors(self, step: int, tensors: Dict[str, np.ndarray], strict: bool = False) -> List[str]:
        missing = []
        with torch.no_grad():
            for name, dst in self.state_tensors().items():
                if name not in tensors:
                    missing.append(name)
                    continue
                src = torch.from_numpy(np.asarray(tensors[name])).to(dst.device, dst.dtype)
                if tuple(src.shape) != tuple(dst.shape):
                    raise ValueError(f"{name}: checkpoint shape {tuple(src.shape)} != model {tuple(dst.shape)}")
                dst.copy_(src)
        hard = [m for m in missing if m in self.variables()]
        if hard and strict:
            raise KeyError(f"variables missing from checkpoint: {hard}")
        self.global_step = int(step)
        if self.backend == "fused":
            self.engine.sync_weights()
            self.engine.set_state(self.opts.batch_size, train=True, step=self.global_step)
        else:
            self.optimizer.step = self.global_step
        return missing

    def export_model(self) -> DeepFM:
        """A CPU DeepFM module holding the current weights (used by export and by tests)."""
        if self.backend == "fused":
            torch.cuda.synchronize(self.device)
            m = DeepFM(self.cfg, seed=0)
            m.load_tf_variables({k: v.detach().cpu() for k, v in self.engine.variables().items()})
            return m
        m = DeepFM(self.cfg, seed=0)
        m.load_tf_variables({k: v.detach().cpu() for k, v in self.model.tf_variables().items()})
        return m

    def launches_per_step(self) -> int:
        return self.engine.launches_per_step() if self.backend == "fused" else 0

    # ------------------------------------------------------------------ training
    def train_batch(self, hb: HostBatch) -> None:
        """One optimizer step on a host batch of any size <= batch_size (the general entry point)."""
        if self.backend == "fused":
            if hb.count == self.opts.batch_size and self.uses_graph and hb.slab is not None and hb.slab.is_pinned():
                self.train_step_host(hb.slab)
                return
            # blocking copies: the batch may be a view of a loader ring slot that is recycled as soon as we return
            ids = hb.feat_ids.to(self.device)
            if ids.dtype != torch.int32:
                ids = ids.int()
            vals = hb.feat_vals.to(self.device)
            labels = hb.labels.to(self.device)
            self.engine.train_step(ids.contiguous(), vals.contiguous(), labels.contiguous())
            self._after_fused_step(count=hb.count)
            if self.uses_graph and hb.count != self.opts.batch_size:
                # the captured graph reads the row count from device state: put the full-batch count back
                self.engine.set_state(self.opts.batch_size, train=True)
        else:
            self._eager_step(hb.feat_ids.to(self.device), hb.feat_vals.to(self.device), hb.labels.to(self.device))

    def train_batches(self, loader, limit: Optional[int] = None):
        """Generator: one optimizer step per batch of ``loader``; yields the batch after its step is queued.

        With the fused back-end and a BatchLoader this is the pipelined path: batch k+1 is fetched before step k is
        launched and handed to the step graph's prefetch node, and a ring slot goes back to the loader only when
        the GPU work that reads it has finished (the host may run ``num_slots - 2`` steps ahead)."""
        from collections import deque
        B = self.opts.batch_size
        deferred = self.backend == "fused" and self.uses_graph and hasattr(loader, "defer_release")
        pinned = {}                # ring slot -> its slab is page-locked (asked once per slot: the query is a driver call)

        def full(hb):
            if hb is None or hb.count != B or hb.slab is None:
                return False
            key = hb.slot if hb.slot is not None else hb.slab.data_ptr()
            if key not in pinned:
                pinned[key] = hb.slab.is_pinned()
            return pinned[key]

        pending = deque()          # (event recorded behind the step that read the slot, slot)
        spare = []                 # events whose step has finished: recorded again instead of creating new ones
        held = set()               # every ring slot not yet given back
        it = iter(loader)

        def fetch():
            hb = next(it, None)
            if hb is not None and deferred:
                held.add(hb.slot)
            return hb

        def give_back(slot):
            if slot in held:
                held.discard(slot)
                loader.release(slot)

        done = 0
        if not deferred:
            # No look-ahead here: asking the loader for the next batch gives the current batch's ring slot back to the
            # parser threads, so a batch must be consumed (copied to the device, synchronously) before the next is fetched.
            for hb in it:
                if limit is not None and done >= limit:
                    break
                self.train_batch(hb)
                done += 1
                yield hb
            return
        if deferred:
            loader.defer_release = True
            self.engine.reset_pipeline()
        try:
            cur = fetch()
            while cur is not None and (limit is None or done < limit):
                if deferred:
                    # Give back every slot whose step has finished (a query, no wait): when the input side is the
                    # slower one the GPU is never far behind, and the parser threads then have nearly the whole ring
                    # to fill ahead.  Wait only when the host has run so far ahead of the GPU that the parser would
                    # be left without a free slot.  (Holding slots until that limit was reached kept the parser to
                    # ONE free slot at all times: 4 parser threads were 56 % busy and the pipeline ran at 19.6 M
                    # examples/s where the same threads decode 33 M/s.)
                    while pending and (len(pending) > max(loader.num_slots - 3, 0) or pending[0][0].query()):
                        ev, slot = pending.popleft()
                        ev.synchronize()
                        give_back(slot)
                        spare.append(ev)
                nxt = fetch() if (limit is None or done + 1 < limit) else None
                if deferred and full(cur):
                    self.train_step_host(cur.slab, nxt.slab if full(nxt) else None)
                else:
                    self.train_batch(cur)
                if deferred:
                    ev = spare.pop() if spare else torch.cuda.Event()
                    ev.record(self.engine.current_stream())
                    pending.append((ev, cur.slot))
                done += 1
                yield cur
                cur = nxt
        finally:
            if deferred:
                torch.cuda.current_stream(self.device).synchronize()
                for slot in list(held):
                    give_back(slot)
                loader.defer_release = False
                self.engine.reset_pipeline()

    def _after_fused_step(self, graph_slot: Optional[int] = None, count: Optional[int] = None) -> None:
        i = self._loss_idx
        self._loss_cnt[i] = self.opts.batch_size if count is None else max(int(count), 1)
        if graph_slot is None:
            self._loss_ring[i].copy_(self.engine.loss_sum_tensor(), non_blocking=True)
        self._loss_src[i] = graph_slot           # None: the trainer's own ring; else: slot of the engine's ring
        self._loss_ev[i].record(self.engine.current_stream())
        self._loss_idx ^= 1
        self._issued += 1
        self.global_step += 1

    def train_step_host(self, host_slab: torch.Tensor, next_slab: Optional[torch.Tensor] = None) -> None:
        """Full batch from a pinned host slab ``[ids | vals | labels]`` (see data.loader.alloc_slab).

        ``next_slab`` (optional): the batch of the following call.  It is copied to the device by a node of this
        step's graph, beside the kernels, so no copy and no cross-stream event sits between two graph launches;
        it must stay unchanged until this step has finished.  The loss must be read (``last_loss``) at least
        every other step for its value to be that of the expected step (the graph writes a 2-entry ring)."""
        tracing.push("train_step")
        if self.backend == "fused":
            if self.uses_graph:
                self._after_fused_step(self.engine.step_pipelined(host_slab, next_slab))
            else:
                ids, vals, labels = self._slab_views(host_slab.to(self.device, non_blocking=True))
                self.engine.train_step(ids, vals, labels)
                self._after_fused_step()
        else:
            ids, vals, labels = self._slab_views(host_slab.to(self.device, non_blocking=True))
            self._eager_step(ids, vals, labels)
        tracing.pop()

    def train_step_device(self, dev_slab: torch.Tensor, next_slab: Optional[torch.Tensor] = None) -> None:
        """Full batch already on the device (same slab layout, same ``next_slab`` contract)."""
        if self.backend == "fused":
            if self.uses_graph:
                self._after_fused_step(self.engine.step_pipelined(dev_slab, next_slab))
            else:
                ids, vals, labels = self._slab_views(dev_slab)
                self.engine.train_step(ids, vals, labels)
                self._after_fused_step()
        else:
            ids, vals, labels = self._slab_views(dev_slab)
            self._eager_step(ids, vals, labels)

    def _slab_views(self, slab: torch.Tensor):
        B, Fd = self.opts.batch_size, self.cfg.field_size
        _, ov, ol = slab_layout(B, Fd, torch.int32)
        return (slab[:ov].view(torch.int32).view(B, Fd), slab[ov:ol].view(torch.float32).view(B, Fd),
                slab[ol:].view(torch.float32))

    def last_loss(self, wait_current: bool = False) -> float:
        """Mean data loss of the previous step (or, with ``wait_current``, of the step just issued).
        Reading one step behind keeps the host from draining the GPU queue."""
        if self.backend == "fused":
            if self._issued == 0:
                return float("nan")
            i = (self._loss_idx ^ 1) if wait_current or self._issued < 2 else self._loss_idx
            self._loss_ev[i].synchronize()
            src = self._loss_src[i]
            total = float(self._loss_ring[i][0]) if src is None else float(self._eng_loss[src])
            return total / self._loss_cnt[i]
        t = self._last_loss_t if wait_current or self._prev_loss_t is None else self._prev_loss_t
        return float(t) if t is not None else float("nan")

    # ---- eager back-end: autograd + restated Horovod collectives ----
    def _eager_step(self, ids: torch.Tensor, vals: torch.Tensor, labels: torch.Tensor) -> None:
        m, opt, cfg = self.model, self.optimizer, self.cfg
        if self.opts.dropout_stream == "engine" or (self.opts.dropout_stream == "auto" and not ids.is_cuda):
            from ..utils.philox import keep_mask
            step, seed, dev = self.global_step + 1, self.opts.seed, ids.device
            m.dropout_mask_fn = lambda layer, rows, cols: keep_mask(seed, step, layer, rows, cols,
                                                                    cfg.dropout_keep[layer], dev)
        ids = ids.long()
        from ..ops.unique_cpu import unique_inverse
        uniq, inverse = unique_inverse(ids, cfg.feature_size)          # torch.unique on a GPU, a table pass on the host
        w_rows = m.fm_w.detach().index_select(0, uniq).requires_grad_(True)
        v_rows = m.fm_v.detach().index_select(0, uniq).requires_grad_(True)
        dense = [(n, p) for n, p in m.trainable_tf_variables().items() if n not in ("fm_w", "fm_v")]
        for _, p in dense:
            p.grad = None
        logits = m.forward_rows(w_rows, v_rows, inverse, vals, training=True)
        loss = m.data_loss(logits, labels)
        loss.backward()
        gw, gv = w_rows.grad, v_rows.grad
        if self.world > 1:
            from ..parallel.sync_dp import allreduce_mean_, sparse_allgather_mean
            allreduce_mean_([p.grad for _, p in dense])
            uniq, gw, gv = sparse_allgather_mean(uniq, gw, gv, cfg.feature_size)
        with torch.no_grad():
            for name, p in dense:
                opt.apply_dense(name, p.data, p.grad)
            if self.opts.table_update_mode == "lazy":
                opt.apply_rows("fm_w", m.fm_w.data, uniq, gw + cfg.l2_reg * m.fm_w.data.index_select(0, uniq))
                opt.apply_rows("fm_v", m.fm_v.data, uniq, gv + cfg.l2_reg * m.fm_v.data.index_select(0, uniq))
            else:
                dw = cfg.l2_reg * m.fm_w.data
                dv = cfg.l2_reg * m.fm_v.data
                dw[uniq] += gw
                dv[uniq] += gv
                opt.apply_dense("fm_w", m.fm_w.data, dw)
                opt.apply_dense("fm_v", m.fm_v.data, dv)
        opt.finish_step()
        self.global_step += 1
        self._prev_loss_t, self._last_loss_t = self._last_loss_t, loss.detach()

    # ------------------------------------------------------------------ inference
    @torch.no_grad()
    def predict_batch(self, hb: HostBatch, labels: bool = False) -> Tuple[torch.Tensor, Optional[float]]:
        """Probabilities [count] (device tensor, valid until the next call) and, optionally, the summed loss."""
        if self.backend == "fused":
            ids = hb.feat_ids.to(self.device, non_blocking=True)
            if ids.dtype != torch.int32:
                ids = ids.int()
            vals = hb.feat_vals.to(self.device, non_blocking=True)
            lab = hb.labels.to(self.device, non_blocking=True) if labels else None
            prob = self.engine.forward(ids.contiguous(), vals.contiguous(), lab)
            loss_sum = None
            if labels:
                loss_sum = float(self.engine.loss_sum_tensor().item())
            return prob, loss_sum
        self.model.eval()
        logits = self.model(hb.feat_ids.to(self.device), hb.feat_vals.to(self.device), training=False)
        loss_sum = None
        if labels:
            loss_sum = float(torch.nn.functional.binary_cross_entropy_with_logits(
                logits, hb.labels.to(self.device), reduction="sum"))
        return torch.sigmoid(logits), loss_sum

    def evaluate(self, batches: Iterable[HostBatch], auc_mode: str = "tf") -> Dict[str, float]:
        """AUC + mean loss over a stream of batches (model_fn EVAL mode, HVD.py:264-278)."""
        tracing.push("evaluate")
        auc = StreamingAUC(200, keep_all=auc_mode in ("exact", "both"))
        loss_sum, n = 0.0, 0
        gpu_hist = None
        dev_pairs = []
        if self.backend == "fused":
            from .engine import auc_histogram
            gpu_hist = torch.zeros(2 * 201, dtype=torch.int64, device=self.device)
        for hb in batches:
            prob, ls = self.predict_batch(hb, labels=True)
            loss_sum += ls
            n += hb.count
            if gpu_hist is not None:
                lab_d = hb.labels.to(self.device)        # blocking: the ring slot is recycled at the next fetch
                auc_histogram(prob, lab_d, gpu_hist)
                if auc.keep_all_enabled:               # exact AUC: predictions stay on the device, sorted there
                    dev_pairs.append((lab_d.clone(), prob.clone()))
            else:
                auc.update(hb.labels.numpy(), prob.cpu().numpy())
        if gpu_hist is not None:
            auc.add_hist(gpu_hist.cpu().numpy())
        reg = 0.0
        v = self.variables()
        l2 = self.cfg.l2_reg
        reg = float(l2 * 0.5 * (v["fm_w"].detach().float().pow(2).sum() + v["fm_v"].detach().float().pow(2).sum()))
        out = {"loss": loss_sum / max(n, 1) + reg, "data_loss": loss_sum / max(n, 1), "examples": n,
               "global_step": self.global_step}
        if self.backend == "fused" and self.uses_graph:
            # forward() changed the device-side count/train flags the captured graph relies on
            self.engine.set_state(self.opts.batch_size, train=True)
        if auc_mode in ("tf", "both"):
            out["auc"] = auc.result()
        if auc_mode in ("exact", "both"):
            if dev_pairs:
                from ..utils.metrics import exact_auc_torch
                out["auc_exact"] = exact_auc_torch(torch.cat([a for a, _ in dev_pairs]), torch.cat([b for _, b in dev_pairs]))
            else:
                out["auc_exact"] = auc.result_exact()
            out.setdefault("auc", out["auc_exact"])
        tracing.pop()
        return out

    def predict(self, batches: Iterable[HostBatch]) -> Iterator[np.ndarray]:
        for hb in batches:
            prob, _ = self.predict_batch(hb, labels=False)
            yield prob.detach().cpu().numpy().copy()
        if self.backend == "fused" and self.uses_graph:
            self.engine.set_state(self.opts.batch_size, train=True)

    def close(self) -> None:
        if self.engine is not None:
            torch.cuda.synchronize(self.device)
            self.engine.close()
            self.engine = None
