"""Inter-graph gap probe: the same captured step replayed back to back with no input copy at all,
vs. the double-buffered copy-stream path.  The difference is what the input plumbing costs."""
import sys, torch
sys.path.insert(0, ".")
import bench
from deepfm_b200.models.deepfm import DeepFMConfig
from deepfm_b200.runtime.trainer import Trainer, TrainerOptions
dev = torch.device("cuda", 0)
cfg = DeepFMConfig(field_size=bench.FIELD, feature_size=bench.FEATURE, embedding_size=bench.EMB, deep_layers=bench.LAYERS,
                   dropout_keep=bench.KEEP, l2_reg=bench.L2)
tr = Trainer(cfg, TrainerOptions(batch_size=2048, optimizer="Adam", learning_rate=bench.LR, backend="fused", seed=1,
                                 cuda_graph=True), device=dev)
pool, _ = bench.make_host_pool(2048, 16, 0, 1)
dpool = [s[0].to(dev) for s in pool]
def timed(fn, n=2000):
    for i in range(50): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
tr.train_step_device(dpool[0]); tr.train_step_device(dpool[1]); torch.cuda.synchronize()
print("replay(0) only, no copy      : %.1f us" % timed(lambda i: tr.engine.replay(0)))
print("replay(i&1), no copy         : %.1f us" % timed(lambda i: tr.engine.replay(i & 1)))
s0 = tr.engine.device_slab(0)[0]
def d2d(i):
    s0.copy_(dpool[i % 16], non_blocking=True); tr.engine.replay(0)
print("in-stream D2D copy + replay  : %.1f us" % timed(d2d))
print("graph prefetch node, device   : %.1f us" % timed(lambda i: tr.train_step_device(dpool[i % 16], dpool[(i + 1) % 16])))
print("graph prefetch node, pinned   : %.1f us" % timed(lambda i: tr.train_step_host(pool[i % 16][0], pool[(i + 1) % 16][0])))
def both(i):
    tr.train_step_host(pool[i % 16][0], pool[(i + 1) % 16][0]); tr.last_loss()
print("  + loss read one step behind : %.1f us" % timed(both))
print("copy-stream double buffering : %.1f us" % timed(lambda i: tr.engine.step_from_host(dpool[i % 16], None)))
print("pinned host slab, e2e        : %.1f us" % timed(lambda i: tr.engine.step_from_host(pool[i % 16][0], None)))
