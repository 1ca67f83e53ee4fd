#!/usr/bin/env python
"""Headline benchmark: DeepFM training throughput (examples/s) on N B200s.

Config (BASELINE.json config 2 / the reference's script defaults, HVD.py:36-50): 39 fields,
feature_size 117,581, embedding 32, MLP 256-128-64, dropout keep 0.5, Adam lr 5e-4 x N, l2 1e-4,
batch 2048 per GPU, synchronous data parallel, replicated embedding table with sparse gradients.
Data: synthetic Criteo-shaped TFRecords written and parsed by the framework's own native IO
layer, random-initialised weights.

  python bench.py --gpus 1 --steps 200 --warmup 20
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
         --master-port 29511 bench.py --gpus 8 --steps 200 --warmup 20

Prints ONE JSON line (rank 0).  `value` times the device work of K full training steps (forward,
backward, optimizer for every parameter) with CUDA events, max over ranks; `e2e.value` times the
same K steps through the public Trainer API including, every step, the pinned-host -> device copy
of that step's batch and a device -> host read of its loss.
`--impl reference` reports that the unmodified TF1/Horovod reference cannot run in this image;
`--impl eager` runs the stock PyTorch (+NCCL) restatement used as the measured baseline.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FIELD, FEATURE, EMB = 39, 117581, 32
LAYERS = (256, 128, 64)
KEEP = (0.5, 0.5, 0.5)
BATCH = 2048
LR, L2 = 5e-4, 1e-4
CONFIGS = {
    "criteo": dict(feature=117581, emb=32, layers=(256, 128, 64),
                   name="DeepFM F39 K32 V117581 MLP256-128-64 dropout0.5 Adam lazy-sparse"),
    "large": dict(feature=33_000_000, emb=64, layers=(400, 400, 400),
                  name="DeepFM F39 K64 V33M MLP400-400-400 dropout0.5 Adam lazy-sparse"),
}
POOL = 256          # distinct batches cycled through: 256 x 647 KB = 166 MB > the 126 MB L2


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--config", default="criteo", choices=sorted(CONFIGS) + ["sweep"],
                    help="criteo = BASELINE.json config 2 (the headline); large = config 3 (33M-row table, emb 64, MLP 400x3); "
                         "sweep = config 5: fused collectives vs NCCL, dense 4 KB-256 MB and sparse 1K-1M ids (needs --gpus >= 2)")
    ap.add_argument("--impl", default="fused", choices=["fused", "reference", "eager", "eager_graph"],
                    help="fused = this framework; reference = the unmodified TF1/Horovod scripts (not installable here); "
                         "eager = the same algorithm on stock PyTorch ops (+NCCL), fp32, launched op by op; eager_graph = "
                         "stock PyTorch at its best: bf16 autocast, fused Adam, NCCL all-reduce, the whole step in one CUDA graph")
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--pool", type=int, default=POOL)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="skip the multi-GPU self-check that precedes the timed region")
    ap.add_argument("--no-pipeline", action="store_true", help="skip the file -> loader -> trainer throughput figure")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.t.join(timeout=1)
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 6:
                continue
            try:
                sm.append(float(p[0])); mx.append(float(p[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def nvlink_kib(index: int):
    """Cumulative NVLink data counters of one GPU (sum over links), KiB: (tx, rx), or None if unavailable."""
    import re
    try:
        out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(index)], capture_output=True, text=True,
                             timeout=20).stdout
    except (OSError, subprocess.TimeoutExpired):
        return None
    tx = sum(int(x) for x in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out))
    rx = sum(int(x) for x in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out))
    return (tx, rx) if (tx or rx) else None


def pool_path(batch: int, pool: int, rank: int) -> str:
    n = batch * pool
    return os.path.join(tempfile.gettempdir(), f"dfm_bench_r{rank}_{n}_{FEATURE}.tfrecords")


def thread_cpu_seconds():
    """CPU seconds (user + system) of every thread of this process, by thread name."""
    out = {}
    try:
        for tid in os.listdir("/proc/self/task"):
            with open(f"/proc/self/task/{tid}/stat") as f:
                st = f.read()
            name = st[st.index("(") + 1:st.rindex(")")]
            f_ = st[st.rindex(")") + 2:].split()
            out[(tid, name)] = (int(f_[11]) + int(f_[12])) / os.sysconf("SC_CLK_TCK")
    except (OSError, ValueError):
        pass
    return out


def measure_pipeline(trainer, path: str, batch: int, seconds: float = 2.5):
    """TFRecord file -> native loader (framing, CRC, Example decode into pinned slabs) -> Trainer.train_batches ->
    loss on the host: what `python -m deepfm_b200.train` achieves per rank, input pipeline included."""
    import torch
    from deepfm_b200.data.loader import BatchLoader
    from deepfm_b200.parallel.sync_dp import chief_decides
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 8)
    workers = max(2, min(32, cores // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))))
    loader = BatchLoader([path], FIELD, batch, num_epochs=100000, drop_remainder=True, num_workers=workers, num_slots=8,
                         ids_dtype=torch.int32, verify_crc=True)
    n, t0, loss, cpu0 = 0, None, float("nan"), {}
    gen = trainer.train_batches(loader)
    try:
        for _ in gen:
            if t0 is None:                       # the first batch pays thread start-up; time from the second on
                torch.cuda.synchronize()
                t0 = time.time()
                cpu0 = thread_cpu_seconds()
                continue
            n += 1
            if n % 64 == 0:
                loss = trainer.last_loss()
            # every rank must launch the same number of steps (each contains collective kernels that wait for the
            # peers' same step), so rank 0's clock decides for all -- asked rarely: the broadcast drains the stream
            if n % 256 == 0 and chief_decides(time.time() - t0 >= seconds, trainer.device):
                break
    finally:
        cpu1, t1 = thread_cpu_seconds(), time.time()
        gen.close()                              # gives the ring slots back before the loader goes away
        loader.close()
    torch.cuda.synchronize()
    dt = time.time() - t0
    # where the host CPU went, in cores (CPU seconds per second of the window), by thread name
    me = str(os.getpid())
    use = {}
    for key, v in cpu1.items():
        name = "main" if key[0] == me else key[1]
        use[name] = use.get(name, 0.0) + (v - cpu0.get(key, 0.0)) / max(t1 - t0, 1e-9)
    cpu = {k: round(v, 2) for k, v in sorted(use.items(), key=lambda kv: -kv[1]) if v >= 0.01}
    return n * batch / dt, dt, workers, loss, cpu, cores


def make_host_pool(batch: int, pool: int, rank: int, world: int):
    """Synthetic Criteo-shaped TFRecords -> native loader -> (list of pinned batch slabs, records/s of the loader alone)."""
    import torch
    pin = torch.cuda.is_available()
    from deepfm_b200.data.loader import BatchLoader, alloc_slab
    from deepfm_b200.data.tfrecord import generate_synthetic
    n = batch * pool
    path = pool_path(batch, pool, rank)
    if not os.path.exists(path):
        generate_synthetic(path + ".tmp", n, FIELD, FEATURE, n_dense=13, seed=1234 + rank, pos_rate=0.2553, zipf_s=1.05)
        os.replace(path + ".tmp", path)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 8)
    workers = max(2, min(32, cores // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))))
    # the loader alone -- framing, CRC32C, Example decode into its ring of slabs, nothing copied out -- four passes
    loader = BatchLoader([path], FIELD, batch, num_epochs=4, drop_remainder=True, num_workers=workers, num_slots=8,
                         ids_dtype=torch.int32, verify_crc=True, pin_memory=pin)
    t0, got = time.time(), 0
    for hb in loader:
        got += hb.count
    rate = got / max(time.time() - t0, 1e-9)
    loader.close()
    # one more pass fills the pool of pinned slabs the timed loops draw from
    loader = BatchLoader([path], FIELD, batch, num_epochs=1, drop_remainder=True, num_workers=workers, num_slots=4,
                         ids_dtype=torch.int32, verify_crc=True, pin_memory=pin)
    slabs = [alloc_slab(batch, FIELD, torch.int32, pin_memory=pin) for _ in range(pool)]
    for i, hb in enumerate(loader):
        slabs[i][0].copy_(hb.slab)
    loader.close()
    return slabs, rate


def run_eager_graph(args):
    """A second, stronger baseline on stock PyTorch: bf16 autocast, torch.optim.Adam(fused, capturable), gradients
    all-reduced with NCCL inside the captured step, the whole step replayed as ONE CUDA graph.  Like the reference it
    treats the table gradient as dense (l2 over the whole table, every row updated every step: HVD.py:266-267)."""
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    world, rank, local_rank = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    if world > 1:
        # the NCCL all-reduce captured inside the step graph did not complete on this box (the run sat in the first replay
        # until it was killed), so this arm is single-GPU; the op-by-op `eager` arm is the multi-GPU stock baseline
        os.dup2(real_stdout, 1)
        if rank == 0:
            print(json.dumps({"impl": "eager_graph", "unavailable": "single-GPU only: NCCL all-reduce inside the captured "
                              "step hung here; use --impl eager for N > 1"}), flush=True)
        return 0
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    cfg = DeepFMConfig(field_size=FIELD, feature_size=FEATURE, embedding_size=EMB, deep_layers=LAYERS, dropout_keep=KEEP, l2_reg=L2)
    model = DeepFM(cfg, seed=1).to(dev)
    params = [p for p in model.parameters()]
    opt = torch.optim.Adam(params, lr=LR * world, fused=True, capturable=True)
    B = args.batch
    pool, parse_rate = make_host_pool(B, args.pool, rank, world)
    P = len(pool)
    ids = torch.zeros(B, FIELD, dtype=torch.int32, device=dev)
    vals = torch.zeros(B, FIELD, device=dev)
    labels = torch.zeros(B, device=dev)
    loss_out = torch.zeros((), device=dev)

    def step():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = model(ids, vals, training=True)
        loss = model.loss(logits.float(), labels)
        loss.backward()
        if world > 1:
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            dist.all_reduce(flat)
            flat.div_(world)
            o = 0
            for p in params:
                p.grad.copy_(flat[o:o + p.numel()].view_as(p.grad))
                o += p.numel()
        opt.step()
        opt.zero_grad(set_to_none=False)
        loss_out.copy_(loss.detach())

    def load(i, non_blocking=True):
        _, hi, hv, hl = pool[i % P]
        ids.copy_(hi, non_blocking=non_blocking); vals.copy_(hv, non_blocking=non_blocking); labels.copy_(hl, non_blocking=non_blocking)

    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        for i in range(3):
            load(i)
            step()
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize(dev)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    K, W = args.steps, max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    dev_pool = [tuple(t.to(dev) for t in pool[i][1:]) for i in range(P)]
    def load_dev(i):
        di, dv, dl = dev_pool[i % P]
        ids.copy_(di); vals.copy_(dv); labels.copy_(dl)
    for i in range(W):
        load_dev(i); graph.replay()
    barrier()
    sampler = ClockSampler(local_rank); sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(K):
        load_dev(W + i); graph.replay()
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms_dev = e0.elapsed_time(e1)
    # end to end: pinned host -> device every step, loss back to the host every step
    host_loss = torch.zeros((), pin_memory=True)
    barrier()
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0.record()
    for i in range(K):
        load(W + i); graph.replay()
        host_loss.copy_(loss_out, non_blocking=True)
    h1.record()
    barrier()
    ms_e2e = h0.elapsed_time(h1)
    t = torch.tensor([ms_dev, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_dev, ms_e2e = t.tolist()
    gb = B * world
    out = {"metric": "train_examples_per_sec", "value": gb * K / (ms_dev / 1e3), "unit": "examples/s", "n_gpus": world,
           "steps": K, "warmup": W, "ms_per_step": ms_dev / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "bf16 autocast", "data": "synthetic", "impl": "eager_graph",
           "config": {"model": CONFIGS[args.config]["name"].replace("lazy-sparse", "dense table update (reference math)"),
                      "global_batch": gb, "seq_len": FIELD, "parallelism": f"dp{world}", "backend": "stock PyTorch, CUDA graph"},
           "clocks": clocks,
           "e2e": {"value": gb * K / (ms_e2e / 1e3), "unit": "examples/s", "ms_per_step": ms_e2e / K,
                   "h2d_bytes_per_step": pool[0][0].numel(), "d2h_bytes_per_step": 4},
           "gpu_launches": 0, "final_loss": float(loss_out)}
    if world > 1:
        dist.destroy_process_group()
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    if rank == 0:
        print(json.dumps(out), flush=True)
    return 0


def run_sweep(args):
    """BASELINE.json config 5: gradient collective sweep, fused NVLink kernels vs torch.distributed (NCCL)."""
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    world, rank, local_rank = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    if world < 2:
        raise SystemExit("--config sweep needs a torchrun launch with >= 2 ranks")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev)
    from deepfm_b200.parallel.collectives import sweep_dense, sweep_sparse
    dense = sweep_dense(dev, [1 << k for k in range(12, 29)])                     # 4 KB .. 256 MB
    sparse = sweep_sparse(dev, [1_000, 4_000, 16_000, 64_000, 256_000, 1_000_000])
    out = {"metric": "collective_sweep", "n_gpus": world, "impl": "fused vs nccl", "unit": "us (device, max over ranks)",
           "dense_allreduce": dense, "sparse_exchange_update": sparse}
    dist.destroy_process_group()
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    if rank == 0:
        print(json.dumps(out), flush=True)
    return 0


def main():
    global FEATURE, EMB, LAYERS
    args = parse_args()
    if args.config == "sweep":
        return run_sweep(args)
    FEATURE, EMB, LAYERS = CONFIGS[args.config]["feature"], CONFIGS[args.config]["emb"], CONFIGS[args.config]["layers"]
    if args.impl == "eager_graph":
        return run_eager_graph(args)
    if args.impl == "reference":
        print(json.dumps({"impl": "reference", "unavailable":
                          "reference is two TF1.14/1.15 + Horovod + sagemaker_tensorflow scripts with no setup.py; "
                          "TensorFlow 1.x has no wheel for Python 3.12/CUDA 12.9 and none is in /opt/wheelhouse "
                          "(see DESIGN.md, BASELINE.md section 2)"}))
        return 0

    # stdout carries exactly one JSON line: while the benchmark runs, file descriptor 1 points at stderr, so what C
    # libraries print (NCCL's version banner when NCCL_DEBUG is set on the box) cannot land in front of it
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit(f"--gpus {args.gpus} needs a torchrun launch with {args.gpus} processes")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from deepfm_b200.models.deepfm import DeepFMConfig
    from deepfm_b200.runtime.trainer import Trainer, TrainerOptions

    check = None
    if world > 1 and args.impl == "fused" and not args.no_check:
        # correctness of the fused collectives on THIS box, before anything is timed: 4 GD steps on fixed per-rank
        # batches, fused NVLink kernels vs eager fp32 + NCCL
        from deepfm_b200.parallel.selfcheck import fused_vs_nccl
        check = fused_vs_nccl(dev, FIELD, FEATURE, EMB, LAYERS, batch=args.batch)
        if not (check["replicas_identical"] and check["worst_rel_update_err_vs_nccl"] < 0.05):
            if rank == 0:
                print("multi-GPU self-check FAILED: " + json.dumps(check), file=sys.stderr, flush=True)
            sys.stdout.flush()
            os.dup2(real_stdout, 1)
            return 3

    cfg = DeepFMConfig(field_size=FIELD, feature_size=FEATURE, embedding_size=EMB, deep_layers=LAYERS,
                       dropout_keep=KEEP, l2_reg=L2)
    opts = TrainerOptions(batch_size=args.batch, optimizer="Adam", learning_rate=LR, table_update_mode="lazy",
                          backend="fused" if args.impl == "fused" else "eager", seed=1,
                          cuda_graph=not args.no_graph, scale_lr_by_world=True)
    trainer = Trainer(cfg, opts, device=dev)

    pool, parse_rate = make_host_pool(args.batch, args.pool, rank, world)
    K, W = args.steps, max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-timed: inputs rotate through a device-resident pool ----------------
    dev_pool = [s[0].to(dev, non_blocking=True) for s in pool]
    torch.cuda.synchronize()
    P = len(dev_pool)
    for i in range(W):
        trainer.train_step_device(dev_pool[i % P], dev_pool[(i + 1) % P])
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    nv0 = nvlink_kib(local_rank) if world > 1 else None
    cnt0 = (trainer.engine.buffer("comm_dbg", dtype="i4").clone().cpu().view(torch.int64)[20:23].tolist()
            if world > 1 and trainer.backend == "fused" else None)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(K):     # every step copies its batch from the pool (inside the step graph, beside the kernels)
        trainer.train_step_device(dev_pool[(W + i) % P], dev_pool[(W + i + 1) % P])
    e1.record()
    barrier()
    clocks = sampler.stop()
    nv1 = nvlink_kib(local_rank) if nv0 else None
    cnt1 = (trainer.engine.buffer("comm_dbg", dtype="i4").clone().cpu().view(torch.int64)[20:23].tolist()
            if cnt0 is not None else None)
    ms_dev = e0.elapsed_time(e1)
    del dev_pool
    if os.environ.get("DFM_BENCH_TIMELINE") and trainer.backend == "fused" and rank == 0:
        # %globaltimer at the start of CTA 0 of every kernel of the last step (StepState::kstamp)
        ks = trainer.engine.buffer("state", dtype="i4").clone().cpu()[12:60].view(torch.int64)
        names = {0: "embed_fwd", 1: "dedupe", 2: "fwd0", 3: "fwd1", 4: "fwd2", 8: "loss", 11: "bwd2", 10: "bwd1",
                 15: "embed_bwd", 16: "wgrad", 17: "dense_opt", 22: "dense_opt_end", 18: "sparse|exchange",
                 19: "sparse_end(cta0)"}
        line = sorted(((int(ks[k]) - int(ks[0])) / 1e3, n) for k, n in names.items() if int(ks[k]) >= int(ks[0]))
        print("timeline us: " + "  ".join(f"{n}@{t:.1f}" for t, n in line), file=sys.stderr, flush=True)
        kd = trainer.engine.buffer("kern_dbg", dtype="i4").clone().cpu().view(torch.int64)
        ph = ["start", "setup", "mma_issued", "staged", "acc_ready", "epilogue", "end"]
        print("embed_bwd CTA(0,0) us: " + "  ".join(f"{ph[i]}@{(int(kd[i]) - int(kd[0])) / 1e3:.1f}" for i in range(7)),
              file=sys.stderr, flush=True)
        n_cta = 80
        st = [int(kd[8 + 2 * i]) for i in range(n_cta)]
        en = [int(kd[9 + 2 * i]) for i in range(n_cta)]
        t0 = min(s for s in st if s > 0)
        print("embed_bwd per-CTA end us: " + " ".join(f"{(e - t0) / 1e3:.0f}" for e in en if e > 0), file=sys.stderr, flush=True)
    if os.environ.get("DFM_BENCH_PHASES") and world > 1 and trainer.backend == "fused":
        # %globaltimer stamps CTA 0 of the exchange kernel left during the last step (ns since its start)
        dbg = trainer.engine.buffer("comm_dbg", dtype="i4").clone().cpu().view(torch.int64)
        print(f"[rank {rank}] exchange phases us (push, barrier1, index, grid barrier, merge+publish, barrier2): " +
              " ".join(f"{(int(dbg[i]) - int(dbg[0])) / 1e3:.1f}" for i in range(1, 7)) +
              " | dense (start vs exchange start; local sum, barrier A, reduce+publish, barrier B, optimizer): " +
              f"{(int(dbg[8]) - int(dbg[0])) / 1e3:.1f}; " +
              " ".join(f"{(int(dbg[i]) - int(dbg[8])) / 1e3:.1f}" for i in (12, 9, 13, 10, 11)), file=sys.stderr, flush=True)

    # ---------------- end to end: pinned host -> device every step, loss back to host ----------------
    for i in range(W):
        trainer.train_step_host(pool[i % P][0], pool[(i + 1) % P][0])
    trainer.last_loss()
    barrier()
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    h0.record()
    loss_acc = 0.0
    for i in range(K):
        # H2D of batch i+1 (pinned -> device) is a node of step i's graph; step 0's own batch was staged by warm-up
        trainer.train_step_host(pool[(W + i) % P][0], pool[(W + i + 1) % P][0])
        loss_acc += trainer.last_loss()          # D2H read of the previous step's loss (pipelined by one step)
    final_loss = trainer.last_loss(wait_current=True)
    h1.record()
    barrier()
    ms_e2e = h0.elapsed_time(h1)

    # ---------------- input pipeline included: TFRecord file -> loader -> train_batches ----------------
    pipe = None
    if not args.no_pipeline:
        rate, dt, workers, _, cpu_use, cores = measure_pipeline(trainer, pool_path(args.batch, args.pool, rank), args.batch)
        pr = torch.tensor([rate], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(pr, op=dist.ReduceOp.MIN)          # the job runs at the pace of its slowest rank
        pipe = {"value": float(pr.item()) * world, "unit": "examples/s", "seconds": round(dt, 2),
                "parser_threads_per_rank": workers, "host_cores": os.cpu_count(), "cores_this_rank_may_use": cores,
                "rank0_cpu_cores_used_by_thread": cpu_use,
                "what": "TFRecord file (page cache) -> framing + CRC32C + Example decode -> pinned slabs -> H2D -> step -> loss"}

    t = torch.tensor([ms_dev, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_dev, ms_e2e = t.tolist()
    global_batch = args.batch * world
    value = global_batch * K / (ms_dev / 1e3)
    e2e = global_batch * K / (ms_e2e / 1e3)
    slab_bytes = pool[0][0].numel()
    out = {
        "metric": "train_examples_per_sec", "value": value, "unit": "examples/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": ms_dev / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16" if args.impl == "fused" else "fp32", "data": "synthetic",
        "impl": args.impl,
        "config": {"model": CONFIGS[args.config]["name"],
                   "global_batch": global_batch, "seq_len": FIELD, "parallelism": f"dp{world}",
                   "l2": f"inputs rotate through {len(pool)} distinct batches ({len(pool) * slab_bytes / 1e6:.0f} MB"
                         + (" > 126 MB L2)" if len(pool) * slab_bytes > 126e6 else
                            "; the embedding tables touched per step exceed L2)" if FEATURE * EMB * 12 > 126e6 else ")"),
                   "backend": trainer.backend, "cuda_graph": trainer.uses_graph},
        "clocks": clocks,
        "e2e": {"value": e2e, "unit": "examples/s", "ms_per_step": ms_e2e / K, "h2d_bytes_per_step": slab_bytes,
                "d2h_bytes_per_step": 4},
        "gpu_launches": trainer.launches_per_step() * K,
        "final_loss": final_loss, "mean_loss": loss_acc / max(K, 1),
        "loader_records_per_sec": parse_rate,
        "pipeline": pipe,
    }
    if check is not None:
        out["replicas_identical"] = check["replicas_identical"]
        out["worst_rel_update_err_vs_nccl"] = check["worst_rel_update_err_vs_nccl"]
        out["comm"] = {"provider": check["provider"], "nvls_dense": bool(check["multicast"] & 1),
                       "nvls_rows": bool(check["multicast"] & 2)}
    if cnt0 is not None:
        # NVLink bytes per step of this rank, from what its kernels COUNTED (records pushed to peers, rows published) --
        # `nvidia-smi nvlink -gt d` reads N/A in this pool's containers, so the link counters cannot be sampled
        steps = max(cnt1[2] - cnt0[2], 1)
        rows, recs = (cnt1[0] - cnt0[0]) / steps, (cnt1[1] - cnt0[1]) / steps
        mc = bool(check and check["multicast"] & 2)
        rec_b, row_b = 4 * EMB + 8, 4 * EMB + 4
        t2 = torch.tensor([rows, recs], device=dev, dtype=torch.float64)
        dist.all_reduce(t2)
        dense_b = 4 * trainer.engine.n_dense * 2.0 * (world - 1) / world       # reduce-scatter + all-gather, each way
        out["nvlink_bytes_per_step"] = {
            "sparse_tx": recs * rec_b + rows * row_b * (1 if mc else world - 1),
            "sparse_rx": (float(t2[1]) / world) * rec_b + (float(t2[0]) - rows) * row_b,
            "dense_each_way": dense_b, "rank": rank, "records_pushed_to_peers": recs, "rows_published": rows,
            "how": "sparse: counted by the exchange kernel (records pushed, rows published; rx from the other ranks' counts); "
                   "dense: size of the vector; the GPUs' link counters read N/A in this pool's containers"}
    if nv0 and nv1:
        # measured by the NVLink counters of this rank's GPU around the device-timed region (KiB granularity)
        out["nvlink_bytes_per_step"] = {"tx": (nv1[0] - nv0[0]) * 1024 / K, "rx": (nv1[1] - nv0[1]) * 1024 / K,
                                        "rank": rank}
    trainer.close()
    if world > 1:
        dist.destroy_process_group()
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    if rank == 0:
        print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
