"""Where the time of one replayed training step goes: every kernel's CTA 0 stamps %globaltimer when it
starts (StepState::kstamp), so the gaps between dependent kernels inside the CUDA graph are visible
without a profiler attached.  Prints the stamps of the last of many back-to-back steps."""
import json, sys
import torch
sys.path.insert(0, ".")
from deepfm_b200.models.deepfm import DeepFMConfig
from deepfm_b200.runtime.trainer import Trainer, TrainerOptions

B = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
cfg = DeepFMConfig(39, 117581, 32, (256, 128, 64), (0.5, 0.5, 0.5), l2_reg=1e-4)
tr = Trainer(cfg, TrainerOptions(batch_size=B, optimizer="Adam", learning_rate=5e-4, backend="fused", seed=1,
                                 cuda_graph=True), device=torch.device("cuda", 0))
g = torch.Generator().manual_seed(0)
from deepfm_b200.data.loader import alloc_slab
pool = []
for i in range(8):
    slab, ids, vals, labels = alloc_slab(B, 39, torch.int32, pin_memory=True)
    u = torch.rand(B, 39, generator=g)
    x = (13 + ((117581 - 600) ** u)).long().clamp(14, 117580).int()
    x[:, :13] = torch.arange(1, 14, dtype=torch.int32)
    ids.copy_(x); v = torch.rand(B, 39, generator=g); v[:, 13:] = 1.0; vals.copy_(v)
    labels.copy_((torch.rand(B, generator=g) < 0.25).float())
    pool.append(slab.cuda())
names = {0: "embed_fwd", 1: "dedupe", 2: "fwd0", 3: "fwd1", 4: "fwd2", 8: "loss", 10: "bwd2", 9 + 1: "bwd1", 15: "embed_bwd",
         16: "wgrad", 17: "dense_opt", 22: "dense_opt_end(cta0)", 18: "sparse", 19: "sparse_end(cta0)"}
names = {0: "embed_fwd", 1: "dedupe", 2: "fwd0", 3: "fwd1", 4: "fwd2", 8: "loss", 11: "bwd(l=2)", 10: "bwd(l=1)",
         15: "embed_bwd", 16: "wgrad", 17: "dense_opt", 22: "dense_opt_end(cta0)", 18: "sparse", 19: "sparse_end(cta0)"}
for rep in range(3):
    for i in range(50):
        tr.train_step_device(pool[i % 8])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(200):
        tr.train_step_device(pool[i % 8])
    e1.record(); torch.cuda.synchronize()
    st = tr.engine.buffer("state", dtype="i4").clone().cpu()
    ks = st[12:12 + 48].view(torch.int64)
    t0 = int(ks[0])
    line = sorted(((int(ks[k]) - t0) / 1e3, n) for k, n in names.items() if int(ks[k]) >= t0)
    print(f"step {e0.elapsed_time(e1) * 5:.1f} us | " + "  ".join(f"{n}@{t:.1f}" for t, n in line), flush=True)
kd = tr.engine.buffer("kern_dbg", dtype="i4").clone().cpu().view(torch.int64)
ph = ["start", "setup", "first_stage", "mma_issued", "acc_ready", "epilogue", "end"]
for name, base in [("fwd0", 384), ("fwd1", 392), ("fwd2", 400), ("bwd(l=2)", 384 + 8 * 8), ("bwd(l=1)", 384 + 8 * 7)]:
    t = [int(kd[base + i]) for i in range(7)]
    print(f"{name:9s} CTA(0,0) us: " + "  ".join(f"{ph[i]}@{(t[i] - t[0]) / 1e3:.2f}" for i in range(7)), flush=True)
import os
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"kstamp": [int(x) for x in ks], "kern_dbg": [int(x) for x in kd]}, open("gpurun_out/timeline_raw.json", "w"))
