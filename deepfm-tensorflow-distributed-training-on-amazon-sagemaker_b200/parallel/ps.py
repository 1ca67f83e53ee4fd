"""Asynchronous parameter-server training on CPU over gloo (the reference's 1-ps-cpu mode).

What TensorFlow's distributed runtime does for ``PS.py`` (``tf.estimator.train_and_evaluate`` with
``TF_CONFIG``; PS.py:461-521, README.md:15-39):

* parameter servers hold the variables -- placed round-robin *per variable* over the PS tasks, as
  ``replica_device_setter`` does (so ``fm_v`` lives whole on one PS) -- and run the optimizer's apply
  ops as gradients arrive, with no locking between workers (Hogwild);
* every worker runs its own input pipeline and model replica, with no synchronisation between
  workers, and the learning rate is NOT scaled by the worker count (PS.py:180);
* the master (worker 0 here) initialises or restores the parameters, the others join once that is
  done (README.md:63), and it alone checkpoints, evaluates on a throttle (PS.py:515-516) and exports.

Re-designed rather than transliterated: a step pulls and pushes only the table rows the batch touches
(the reference moves the whole 15 MB ``fm_v`` both ways every step because of its dense L2 term,
SURVEY X6); the L2 term is applied on the PS to the rows being updated (``lazy``) or to all rows
(``dense_exact``).  Transport is ``torch.distributed`` point-to-point on the gloo backend.

Process layout: ranks ``[0, n_ps)`` are parameter servers, the rest are workers; launch with
``torchrun --nproc-per-node (n_ps + n_workers) -m deepfm_b200.train --dist ps --ps_hosts a,b ...``
(the number of names in ``--ps_hosts`` is the number of PS tasks), or one process per role with
``--job_name ps|worker --task_index i`` plus ``--ps_hosts/--worker_hosts``.
"""
from __future__ import annotations

import os
import time
from collections import OrderedDict
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

from ..models.deepfm import DeepFM, DeepFMConfig
from ..utils.optimizers import TFOptimizer

# request codes (header = int64[4]: op, variable index, n, aux)
OP_PULL_ROWS, OP_PULL, OP_PUSH_ROWS, OP_PUSH, OP_SET, OP_STEP, OP_GET_SLOTS, OP_INIT_DONE, OP_IS_INIT, \
    OP_DONE, OP_N_DONE, OP_ASSIGN, OP_PULL_DENSE, OP_PUSH_DENSE = range(14)

TABLES = ("fm_w", "fm_v")


def variable_placement(names: List[str], n_ps: int) -> Dict[str, int]:
    """Round-robin per variable in creation order (tf.train.replica_device_setter's default strategy)."""
    return {n: i % n_ps for i, n in enumerate(names)}


def _hdr(op: int, var: int = 0, n: int = 0, aux: int = 0) -> torch.Tensor:
    return torch.tensor([op, var, n, aux], dtype=torch.int64)


# =============================================================================== parameter server
class ParameterServer:
    def __init__(self, cfg: DeepFMConfig, rank: int, n_ps: int, n_workers: int, optimizer: str, lr: float,
                 table_update_mode: str = "lazy"):
        self.cfg, self.rank, self.n_ps, self.n_workers = cfg, rank, n_ps, n_workers
        template = DeepFM(cfg, seed=0)
        self.names = list(template.tf_variables().keys())
        self.place = variable_placement(self.names, n_ps)
        # host-resident shard: only the variables placed on this task
        self.vars: "OrderedDict[str, torch.Tensor]" = OrderedDict(
            (n, torch.zeros_like(t.detach())) for n, t in template.tf_variables().items() if self.place[n] == rank)
        self.opt = TFOptimizer(optimizer, lr)
        for n, t in self.vars.items():
            if "moving_" not in n:
                self.opt.ensure_slots(n, t)
        self.mode = table_update_mode
        # everything here that is not an embedding table, in variable order: one message moves all of it
        self.dense_names = [n for n in self.vars if n not in TABLES]
        self.initialised = False
        self.done = 0
        self.global_step = 0
        self.applies = 0

    def serve(self) -> None:
        """Single-threaded request loop; ends when every worker has said it is done."""
        hdr = torch.zeros(4, dtype=torch.int64)
        t_start = t_mark = time.perf_counter()
        self.idle_s, self.op_s = 0.0, {}
        last_op = None
        while self.done < self.n_workers:
            now = time.perf_counter()
            if last_op is not None:                          # time spent serving the previous request, by kind
                self.op_s[last_op] = self.op_s.get(last_op, 0.0) + now - t_mark
            src = dist.recv(hdr)
            t_mark = time.perf_counter()
            self.idle_s += t_mark - now                      # waiting for a request
            op, vi, n, aux = (int(x) for x in hdr)
            last_op = op
            name = self.names[vi]
            if op == OP_PULL_ROWS:
                ids = torch.empty(n, dtype=torch.int64)
                dist.recv(ids, src=src)
                dist.send(self.vars[name][ids].contiguous(), dst=src)
            elif op == OP_PULL:
                dist.send(self.vars[name].contiguous(), dst=src)
            elif op == OP_PUSH_ROWS:
                p = self.vars[name]
                ids = torch.empty(n, dtype=torch.int64)
                g = torch.empty((n,) + tuple(p.shape[1:]), dtype=p.dtype)
                dist.recv(ids, src=src)
                dist.recv(g, src=src)
                l2 = self.cfg.l2_reg
                if self.mode == "lazy":
                    self.opt.apply_rows(name, p, ids, g + l2 * p[ids])
                else:                                   # every row sees l2 * V, as in the reference graph
                    dense = l2 * p
                    dense[ids] += g
                    self.opt.apply_dense(name, p, dense)
                self.applies += 1
            elif op == OP_PUSH:
                g = torch.empty_like(self.vars[name])
                dist.recv(g, src=src)
                self.opt.apply_dense(name, self.vars[name], g)
                self.applies += 1
            elif op == OP_PULL_DENSE:
                # every dense variable (and BN moving statistic) of this task as ONE flat message
                flat = torch.cat([self.vars[k].reshape(-1) for k in self.dense_names]) if self.dense_names \
                    else torch.zeros(0)
                dist.send(flat, dst=src)
            elif op == OP_PUSH_DENSE:
                # one flat message: gradients of the trainable dense variables, new values of the moving statistics
                flat = torch.empty(sum(self.vars[k].numel() for k in self.dense_names))
                dist.recv(flat, src=src)
                o = 0
                for k in self.dense_names:
                    t = self.vars[k]
                    piece = flat[o:o + t.numel()].view_as(t)
                    o += t.numel()
                    if "moving_" in k:
                        t.copy_(piece)                   # assigned, not optimised
                    else:
                        self.opt.apply_dense(k, t, piece)
                self.applies += 1
            elif op in (OP_SET, OP_ASSIGN):
                t = torch.empty_like(self.vars[name])
                dist.recv(t, src=src)
                if aux == 0:
                    self.vars[name].copy_(t)
                else:                                  # aux = 1 + slot index: restoring optimizer slots
                    k = self.opt.slot_names(name)[aux - 1]
                    if k:
                        self.opt.slots[k].copy_(t)
            elif op == OP_STEP:
                # one worker finished one train_op: advance this task's view of the step (Adam's beta powers)
                self.opt.finish_step()
                self.global_step += 1
                dist.send(torch.tensor([self.global_step], dtype=torch.int64), dst=src)
            elif op == OP_GET_SLOTS:
                s0, s1 = self.opt.slot_names(name)
                k = s0 if aux == 0 else s1
                t = self.opt.slots.get(k) if k else None
                dist.send((t if t is not None else torch.zeros_like(self.vars[name])).contiguous(), dst=src)
            elif op == OP_INIT_DONE:
                self.initialised = True
                self.global_step = aux
                self.opt.step = aux
            elif op == OP_IS_INIT:
                dist.send(torch.tensor([int(self.initialised)], dtype=torch.int64), dst=src)
            elif op == OP_DONE:
                self.done += 1
            elif op == OP_N_DONE:
                dist.send(torch.tensor([self.done], dtype=torch.int64), dst=src)
            else:
                raise RuntimeError(f"parameter server {self.rank}: unknown request {op} from rank {src}")
        self.wall_s = time.perf_counter() - t_start
        if os.environ.get("DFM_PS_STATS"):
            names = {OP_PULL_ROWS: "pull_rows", OP_PUSH_ROWS: "push_rows", OP_PULL_DENSE: "pull_dense",
                     OP_PUSH_DENSE: "push_dense", OP_STEP: "step"}
            busy = {names.get(k, str(k)): round(v, 3) for k, v in sorted(self.op_s.items(), key=lambda kv: -kv[1])}
            print(f"[ps {self.rank}] wall {self.wall_s:.2f} s, idle {self.idle_s:.2f} s, busy by request {busy}, "
                  f"variables {list(self.vars)}", flush=True)


# =============================================================================== worker side
class PSClient:
    """What a worker sees: variables by name, wherever they live."""

    def __init__(self, cfg: DeepFMConfig, n_ps: int):
        template = DeepFM(cfg, seed=0)
        self.shapes = OrderedDict((n, tuple(t.shape)) for n, t in template.tf_variables().items())
        self.names = list(self.shapes.keys())
        self.index = {n: i for i, n in enumerate(self.names)}
        self.place = variable_placement(self.names, n_ps)
        self.n_ps = n_ps

    def _req(self, name: str, op: int, n: int = 0, aux: int = 0) -> int:
        dst = self.place[name]
        dist.send(_hdr(op, self.index[name], n, aux), dst=dst)
        return dst

    def pull(self, name: str) -> torch.Tensor:
        dst = self._req(name, OP_PULL)
        t = torch.empty(self.shapes[name], dtype=torch.float32)
        dist.recv(t, src=dst)
        return t

    def pull_rows(self, name: str, ids: torch.Tensor) -> torch.Tensor:
        dst = self._req(name, OP_PULL_ROWS, ids.numel())
        dist.send(ids.contiguous(), dst=dst)
        t = torch.empty((ids.numel(),) + self.shapes[name][1:], dtype=torch.float32)
        dist.recv(t, src=dst)
        return t

    def push(self, name: str, grad: torch.Tensor) -> None:
        dist.send(grad.contiguous(), dst=self._req(name, OP_PUSH))

    def push_rows(self, name: str, ids: torch.Tensor, grad: torch.Tensor) -> None:
        dst = self._req(name, OP_PUSH_ROWS, ids.numel())
        dist.send(ids.contiguous(), dst=dst)
        dist.send(grad.contiguous(), dst=dst)

    def set(self, name: str, value: torch.Tensor, slot: int = 0, assign: bool = False) -> None:
        dist.send(value.detach().float().contiguous(), dst=self._req(name, OP_ASSIGN if assign else OP_SET, aux=slot))

    def get_slot(self, name: str, which: int) -> torch.Tensor:
        dst = self._req(name, OP_GET_SLOTS, aux=which)
        t = torch.empty(self.shapes[name], dtype=torch.float32)
        dist.recv(t, src=dst)
        return t

    def dense_names_on(self, ps: int) -> List[str]:
        return [n for n in self.names if self.place[n] == ps and n not in TABLES]

    def pull_dense_all(self) -> Dict[str, torch.Tensor]:
        """All dense variables and BN statistics: one request + one flat reply per PS task."""
        out: Dict[str, torch.Tensor] = {}
        for ps in range(self.n_ps):
            names = self.dense_names_on(ps)
            if not names:
                continue
            dist.send(_hdr(OP_PULL_DENSE), dst=ps)
            flat = torch.empty(sum(int(np.prod(self.shapes[n])) for n in names), dtype=torch.float32)
            dist.recv(flat, src=ps)
            o = 0
            for n in names:
                k = int(np.prod(self.shapes[n]))
                out[n] = flat[o:o + k].view(self.shapes[n])
                o += k
        return out

    def push_dense_all(self, values: Dict[str, torch.Tensor]) -> None:
        """``values[name]`` = gradient of a trainable dense variable, or the new value of a moving statistic."""
        for ps in range(self.n_ps):
            names = self.dense_names_on(ps)
            if not names:
                continue
            dist.send(_hdr(OP_PUSH_DENSE), dst=ps)
            dist.send(torch.cat([values[n].detach().reshape(-1).float() for n in names]), dst=ps)

    # -- one training step's traffic, requests first and replies after, so the PS tasks serve it concurrently --------
    def pull_step(self, uniq: torch.Tensor) -> Tuple[Dict[str, torch.Tensor], torch.Tensor, torch.Tensor]:
        """(dense variables by name, fm_w rows, fm_v rows) for the distinct ids ``uniq``.  Every request of the step is
        sent before the first reply is awaited: messages between two processes keep their order, a PS answers its
        requests in the order they arrive, so the replies come back in the order asked -- and the tasks work at the
        same time instead of one round trip after another."""
        uniq = uniq.contiguous()
        dense_ps = [ps for ps in range(self.n_ps) if self.dense_names_on(ps)]
        # receives are posted BEFORE anything is sent: a PS blocks in its reply until the worker takes it, and would
        # then never read the worker's next request (both ends stuck in send)
        flats, waits = {}, []
        for ps in dense_ps:
            names = self.dense_names_on(ps)
            flats[ps] = torch.empty(sum(int(np.prod(self.shapes[n])) for n in names), dtype=torch.float32)
        rows = {name: torch.empty((uniq.numel(),) + self.shapes[name][1:], dtype=torch.float32) for name in TABLES}
        # per PS task the replies arrive in request order: dense first, then the tables in TABLES order
        for ps in dense_ps:
            waits.append(dist.irecv(flats[ps], src=ps))
        for name in TABLES:
            waits.append(dist.irecv(rows[name], src=self.place[name]))
        for ps in dense_ps:
            dist.send(_hdr(OP_PULL_DENSE), dst=ps)
        for name in TABLES:
            dist.send(uniq, dst=self._req(name, OP_PULL_ROWS, uniq.numel()))
        for w in waits:
            w.wait()
        out: Dict[str, torch.Tensor] = {}
        for ps in dense_ps:
            o = 0
            for n in self.dense_names_on(ps):
                k = int(np.prod(self.shapes[n]))
                out[n] = flats[ps][o:o + k].view(self.shapes[n])
                o += k
        return out, rows[TABLES[0]], rows[TABLES[1]]

    def push_step(self, uniq: torch.Tensor, gw: torch.Tensor, gv: torch.Tensor, values: Dict[str, torch.Tensor]) -> int:
        """Sparse gradients, dense gradients / moving statistics and the end-of-step notice to every PS task, all
        sent before the step replies are read; returns the global step (kept on task 0)."""
        self.push_rows("fm_w", uniq, gw)
        self.push_rows("fm_v", uniq, gv)
        self.push_dense_all(values)
        steps = [torch.zeros(1, dtype=torch.int64) for _ in range(self.n_ps)]
        waits = [dist.irecv(steps[ps], src=ps) for ps in range(self.n_ps)]
        for ps in range(self.n_ps):
            dist.send(_hdr(OP_STEP), dst=ps)
        for w in waits:
            w.wait()
        return int(steps[0])

    def step_all(self) -> int:
        """Tells every PS task one train_op finished; returns the global step (kept on task 0)."""
        gs = 0
        for ps in range(self.n_ps):
            dist.send(_hdr(OP_STEP), dst=ps)
            t = torch.zeros(1, dtype=torch.int64)
            dist.recv(t, src=ps)
            if ps == 0:
                gs = int(t)
        return gs

    def broadcast_simple(self, op: int, aux: int = 0) -> None:
        for ps in range(self.n_ps):
            dist.send(_hdr(op, 0, 0, aux), dst=ps)

    def ask(self, op: int, ps: int = 0) -> int:
        dist.send(_hdr(op), dst=ps)
        t = torch.zeros(1, dtype=torch.int64)
        dist.recv(t, src=ps)
        return int(t)


class PSWorker:
    """One model replica.  ``train_batch`` = pull touched state, forward/backward locally, push gradients."""

    def __init__(self, cfg: DeepFMConfig, client: PSClient, seed: int = 0, dropout_seed: Optional[int] = None):
        """``dropout_seed``: key of this worker's dropout stream (default ``seed``; give every worker its own).  The
        masks come from the counter-based generator the GPU engine uses (utils/philox.py, native on the host): a
        tenth of the cost of torch's scalar CPU dropout, and reproducible per (seed, local step, layer, row)."""
        self.cfg, self.c = cfg, client
        self.model = DeepFM(cfg, seed=seed)
        self.dropout_seed = int(seed if dropout_seed is None else dropout_seed)
        self.local_step = 0
        self.dense_names = [n for n in client.names if n not in TABLES and "moving_" not in n]
        self.stat_names = [n for n in client.names if "moving_" in n]
        self.global_step = 0
        self.last_loss = float("nan")

    def pull_dense(self) -> None:
        tv = self.model.tf_variables()
        with torch.no_grad():
            for n, t in self.c.pull_dense_all().items():
                tv[n].copy_(t)

    def pull_all(self) -> None:
        tv = self.model.tf_variables()
        with torch.no_grad():
            for n in self.c.names:
                tv[n].copy_(self.c.pull(n))

    def train_batch(self, ids: torch.Tensor, vals: torch.Tensor, labels: torch.Tensor) -> float:
        m = self.model
        self.local_step += 1
        if any(k < 1.0 for k in self.cfg.dropout_keep):
            from ..utils.philox import keep_mask
            seed, step, keeps = self.dropout_seed, self.local_step, self.cfg.dropout_keep
            m.dropout_mask_fn = lambda layer, rows, cols: keep_mask(seed, step, layer, rows, cols, keeps[layer])
        from ..ops.unique_cpu import unique_inverse
        uniq, inverse = unique_inverse(ids, self.cfg.feature_size)
        dense, w_rows, v_rows = self.c.pull_step(uniq)
        tv = m.tf_variables()
        with torch.no_grad():
            for n, t in dense.items():
                tv[n].copy_(t)
        w_rows.requires_grad_(True)
        v_rows.requires_grad_(True)
        for n in self.dense_names:
            tv[n].grad = None
        loss = m.data_loss(m.forward_rows(w_rows, v_rows, inverse, vals, training=True), labels)
        loss.backward()
        # dense gradients and the BN moving averages (assigned, not optimised) travel as one message per PS task
        self.global_step = self.c.push_step(uniq, w_rows.grad, v_rows.grad,
                                            {**{n: tv[n].grad for n in self.dense_names},
                                             **{n: tv[n] for n in self.stat_names}})
        self.last_loss = float(loss.detach())
        return self.last_loss


# =============================================================================== driver
def _cluster(flags, env) -> Tuple[int, int, str, int]:
    """(n_ps, n_workers, role, index) from --ps_hosts/--worker_hosts/--job_name/--task_index, TF_CONFIG,
    or a torchrun world (first n_ps ranks are PS tasks)."""
    import json
    tf_config = os.environ.get("TF_CONFIG")
    if tf_config:                                        # the SageMaker / Estimator convention (PS.py:462)
        tc = json.loads(tf_config)
        cl, task = tc.get("cluster", {}), tc.get("task", {})
        n_ps = len(cl.get("ps", [])) or 1
        n_workers = len(cl.get("master", [])) + len(cl.get("chief", [])) + len(cl.get("worker", []))
        ttype, tidx = task.get("type", "worker"), int(task.get("index", 0))
        if ttype == "ps":
            return n_ps, n_workers, "ps", tidx
        if ttype == "evaluator":                         # outside the training cluster (set_dist_env mode 2)
            return n_ps, n_workers, "evaluator", tidx
        has_chief = bool(cl.get("master") or cl.get("chief"))
        return n_ps, n_workers, "worker", tidx + (1 if ttype == "worker" and has_chief else 0)
    n_ps = len([h for h in str(flags.ps_hosts).split(",") if h]) or int(os.environ.get("DFM_NUM_PS", "1"))
    if flags.job_name:
        n_workers = len([h for h in str(flags.worker_hosts).split(",") if h]) or 1
        role = "ps" if flags.job_name == "ps" else "worker"
        return n_ps, n_workers, role, int(flags.task_index)
    world = env.world_size
    if world <= n_ps:
        raise SystemExit(f"parameter-server mode needs more processes ({world}) than PS tasks ({n_ps})")
    return n_ps, world - n_ps, ("ps" if env.rank < n_ps else "worker"), (env.rank if env.rank < n_ps else env.rank - n_ps)


def run_ps(flags, env) -> int:
    """Entry point of ``--dist ps``: this process becomes a parameter server or a worker."""
    import glob
    from ..data.loader import BatchLoader
    from ..data.sharding import ps_shard
    from ..utils.checkpoint import CheckpointManager
    from ..utils.export import export_servable
    from ..utils.metrics import StreamingAUC

    n_ps, n_workers, role, index = _cluster(flags, env)
    if role == "evaluator":
        print("[ps] evaluator task: evaluation is run by the chief on its throttle (PS.py:501-521); nothing to do")
        return 0
    rank = index if role == "ps" else n_ps + index
    world = n_ps + n_workers
    os.environ.setdefault("MASTER_ADDR", env.master_addr)
    os.environ.setdefault("MASTER_PORT", str(env.master_port))
    # intra-op threading as in PS.py:453-459 (OMP/MKL threads = CPUs of the host, shared by co-located tasks)
    torch.set_num_threads(max(1, int(flags.num_CPUs) // max(1, env.local_world_size)))
    if not dist.is_initialized():
        dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = DeepFMConfig(field_size=flags.field_size, feature_size=flags.feature_size,
                       embedding_size=flags.embedding_size, deep_layers=flags.deep_layer_list,
                       dropout_keep=flags.dropout_keep_list, batch_norm=flags.batch_norm,
                       batch_norm_decay=flags.batch_norm_decay, l2_reg=flags.l2_reg, loss_type=flags.loss_type)
    tag = f"[{role} {index}]"
    if role == "ps":
        ps = ParameterServer(cfg, index, n_ps, n_workers, flags.optimizer, flags.learning_rate, flags.table_update_mode)
        print(tag, f"serving {list(ps.vars.keys())}", flush=True)
        ps.serve()
        print(tag, f"done after {ps.applies} applies, global_step {ps.global_step}", flush=True)
        dist.destroy_process_group()
        return 0

    client = PSClient(cfg, n_ps)
    worker = PSWorker(cfg, client, seed=flags.seed, dropout_seed=flags.seed * 1000003 + index)
    chief = index == 0
    model_dir = flags.effective_model_dir
    ckpt = CheckpointManager(model_dir if chief else "", flags.keep_checkpoint_max, flags.save_checkpoints_secs)

    def snapshot() -> Dict[str, np.ndarray]:
        out = {}
        for n in client.names:
            out[n] = client.pull(n).numpy()
            if "moving_" in n:
                continue
            for which, sname in enumerate(TFOptimizer(flags.optimizer, 1.0).slot_names(n)):
                if sname:
                    out[sname] = client.get_slot(n, which).numpy()
        return out

    if chief:
        # the master initialises (or restores) every variable on its PS task, then opens the gate
        start_step = 0
        restored = CheckpointManager(model_dir).restore_latest() if model_dir else None
        tv = worker.model.tf_variables()
        if restored is not None:
            start_step, tensors = restored
            worker.model.load_tf_variables({k: torch.from_numpy(v) for k, v in tensors.items() if k in tv}, strict=False)
            print(tag, f"Restoring parameters from {model_dir} at step {start_step}", flush=True)
        for n in client.names:
            client.set(n, tv[n])
            if restored is not None and "moving_" not in n:
                for which, sname in enumerate(TFOptimizer(flags.optimizer, 1.0).slot_names(n)):
                    if sname and sname in restored[1]:
                        client.set(n, torch.from_numpy(restored[1][sname]), slot=which + 1)
        client.broadcast_simple(OP_INIT_DONE, aux=start_step)
    else:
        while not client.ask(OP_IS_INIT):                # "others join once the master is up" (README.md:63)
            time.sleep(0.05)

    # SageMaker pipe mode (PS.py:137-151): the channels are FIFOs /opt/ml/input/data/<channel>_<epoch>, read as streams
    pipe = bool(flags.pipe_mode)

    def channel_fifo(name: str, epoch: int = 0) -> str:
        return os.path.join(os.environ.get("SM_INPUT_DATA_DIR", "/opt/ml/input/data"), f"{name}_{epoch}")

    train_channel = flags.training_channel_name or "training"
    eval_channel = flags.evaluation_channel_name or "evaluation"

    def evaluate(files) -> Dict[str, float]:
        worker.pull_all()
        worker.model.eval()
        auc, loss_sum, n = StreamingAUC(200), 0.0, 0
        loader = BatchLoader(files, cfg.field_size, flags.batch_size, drop_remainder=flags.drop_remainder_eval,
                             ids_dtype=torch.int64, num_workers=flags.parser_threads, streaming=pipe, feature_size=flags.feature_size)
        with torch.no_grad():
            for hb in loader:
                logits = worker.model(hb.feat_ids, hb.feat_vals, training=False)
                loss_sum += float(torch.nn.functional.binary_cross_entropy_with_logits(logits, hb.labels, reduction="sum"))
                auc.update(hb.labels.numpy(), torch.sigmoid(logits).numpy())
                n += hb.count
        loader.close()
        return {"auc": auc.result(), "loss": loss_sum / max(n, 1), "examples": n, "global_step": worker.global_step}

    if pipe:
        tr = [channel_fifo(train_channel, e) for e in range(flags.num_epochs)]       # one FIFO per epoch
        va = [channel_fifo(eval_channel, 0)]
        te = [channel_fifo(eval_channel, 0)]
    else:
        tr = sorted(glob.glob(f"{flags.training_data_dir}/**/tr*.tfrecords", recursive=True)) if flags.training_data_dir else []
        va = sorted(glob.glob(f"{flags.val_data_dir}/**/va*.tfrecords", recursive=True)) if flags.val_data_dir else []
        te = sorted(glob.glob(f"{flags.val_data_dir}/**/te*.tfrecords", recursive=True)) if flags.val_data_dir else []
    if flags.task_type == "train":
        if not tr:
            raise SystemExit(f"no tr*.tfrecords under --training_data_dir {flags.training_data_dir!r}")
        shard = ps_shard(flags.enable_s3_shard, flags.hosts, flags.current_host)      # PS.py:153-156
        if shard == (1, 0) and len(flags.hosts) <= 1 and n_workers > 1 and not flags.enable_s3_shard:
            shard = (n_workers, index)       # several workers on one host: split the stream between them
        loader = BatchLoader(tr, cfg.field_size, flags.batch_size, num_epochs=1 if pipe else flags.num_epochs,
                             drop_remainder=True, shard_n=shard[0], shard_i=shard[1], ids_dtype=torch.int64,
                             num_workers=flags.parser_threads, streaming=pipe, feature_size=flags.feature_size)
        t_start = t_log = time.time()
        last_eval, steps = t_start, 0
        for hb in loader:
            loss = worker.train_batch(hb.feat_ids, hb.feat_vals, hb.labels)
            steps += 1
            if flags.log_steps and steps % flags.log_steps == 0:
                dt = max(time.time() - t_log, 1e-9)
                print(tag, f"loss = {loss:.6f}, step = {worker.global_step} ({flags.log_steps / dt:.1f} local steps/sec)", flush=True)
                t_log = time.time()
            if chief:
                if ckpt.due():
                    print(tag, "Saving checkpoints for", worker.global_step, "into", ckpt.save(worker.global_step, snapshot()), flush=True)
                now = time.time()
                if va and now - t_start >= flags.ps_eval_start_delay_secs and now - last_eval >= flags.ps_eval_throttle_secs:
                    print(tag, "eval:", evaluate(va), flush=True)
                    last_eval = time.time()
            if flags.max_steps and steps >= flags.max_steps:
                break
        loader.close()
        if chief:
            # a finished worker never blocks the others (README.md:37); the master waits for them, then does the
            # last save / evaluation / export while the parameter servers are still up, and only then lets go
            while client.ask(OP_N_DONE) < n_workers - 1:
                time.sleep(0.05)
            if ckpt.model_dir:
                print(tag, "Saving checkpoints for", worker.global_step, "into", ckpt.save(worker.global_step, snapshot()), flush=True)
            if va:
                print(tag, "eval:", evaluate(va), flush=True)
            if flags.servable_model_dir:
                worker.pull_all()
                print(tag, "servable model exported to", export_servable(worker.model, flags.servable_model_dir), flush=True)
    elif chief:
        # eval / infer / export run on the chief against the parameters the servers hold (restored above), like
        # Estimator.evaluate / predict / export_savedmodel in PS.py:522-551; the other workers only say goodbye
        if flags.task_type == "eval":
            print(tag, "eval:", evaluate(va), flush=True)
        elif flags.task_type == "infer":                 # PS.py:526-533: one "%f" line per example into pred.txt
            worker.pull_all()
            worker.model.eval()
            out_path = os.path.join(flags.val_data_dir or ".", "pred.txt")
            loader = BatchLoader(te, cfg.field_size, flags.batch_size, drop_remainder=flags.drop_remainder_eval,
                                 ids_dtype=torch.int64, num_workers=flags.parser_threads, streaming=pipe, feature_size=flags.feature_size)
            n = 0
            with open(out_path, "w") as fo, torch.no_grad():
                for hb in loader:
                    prob = torch.sigmoid(worker.model(hb.feat_ids, hb.feat_vals, training=False))
                    for pv in prob.tolist():
                        fo.write("%f\n" % pv)
                    n += hb.count
            loader.close()
            print(tag, f"{n} predictions written to {out_path}", flush=True)
        elif flags.task_type == "export":                # PS.py:535-551
            if not flags.servable_model_dir:
                raise SystemExit("--task_type export needs --servable_model_dir")
            worker.pull_all()
            print(tag, "servable model exported to", export_servable(worker.model, flags.servable_model_dir), flush=True)
    client.broadcast_simple(OP_DONE)       # every worker, whatever the task: lets the parameter servers finish

    dist.destroy_process_group()
    return 0
