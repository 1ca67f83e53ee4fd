"""Warm per-kernel device times of one fused training step (CUPTI activity tracing via torch.profiler;
not ncu: kernels run back to back in their real order with warm caches)."""
import json, sys
import torch
from torch.profiler import profile, ProfilerActivity
sys.path.insert(0, ".")
from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.runtime.engine import FusedEngine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
cfg = DeepFMConfig(39, 117581, 32, (256, 128, 64), (0.5, 0.5, 0.5), l2_reg=1e-4)
eng = FusedEngine(cfg, B, optimizer="Adam", lr=5e-4)
eng.load_from_module(DeepFM(cfg, seed=1))
g = torch.Generator(device="cuda").manual_seed(0)
pool = []
for i in range(8):
    u = torch.rand(B, 39, device="cuda", generator=g)
    ids = (13 + ((117581 - 600) ** u)).long().clamp(14, 117580)
    ids[:, :13] = torch.arange(1, 14, device="cuda")
    vals = torch.rand(B, 39, device="cuda", generator=g); vals[:, 13:] = 1.0
    labels = (torch.rand(B, device="cuda", generator=g) < 0.25).float()
    pool.append((ids.int().contiguous(), vals, labels))
eng.set_state(B, train=True)
for i in range(10):
    eng.train_step(*pool[i % 8], set_state=False)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(40):
        eng.train_step(*pool[i % 8], set_state=False)
    torch.cuda.synchronize()
rows = sorted(((e.key[:70], e.device_time_total / max(e.count, 1), e.count) for e in prof.key_averages()
               if e.device_time_total > 0), key=lambda r: -r[1] * r[2])
tot = 0.0
for k, t, c in rows:
    per_step = t * c / 40
    tot += per_step
    print(f"{k:70s} {t:8.2f} us x{c / 40:.0f}/step = {per_step:7.2f}")
print(f"sum of kernel time per step: {tot:.1f} us (two streams overlap part of it)")

dbg = eng.buffer("kern_dbg", dtype="i4").clone().cpu().view(torch.int64)
names = ["start", "setup_done", "mma_issued", "staging_done", "acc_ready", "epilogue_done", "end"]
print("embed_bwd CTA(0,0) phases (us):", {n: round((int(dbg[i]) - int(dbg[0])) / 1000.0, 2) for i, n in enumerate(names)})

t0 = min(int(dbg[8 + 2 * c]) for c in range(80))
rows = []
for y in range(5):
    st = [round((int(dbg[8 + 2 * (y * 16 + x)]) - t0) / 1000.0, 1) for x in range(16)]
    en = [round((int(dbg[9 + 2 * (y * 16 + x)]) - t0) / 1000.0, 1) for x in range(16)]
    print(f"slab {y}: start {min(st)}..{max(st)}  end {min(en)}..{max(en)}  dur {min(e - s for s, e in zip(st, en)):.1f}..{max(e - s for s, e in zip(st, en)):.1f}")

for x in range(16):
    base = 200 + x * 8
    print(f"slab3 tile {x:2d}:", {n: round((int(dbg[base + i]) - int(dbg[base])) / 1000.0, 1) for i, n in enumerate(names)})
