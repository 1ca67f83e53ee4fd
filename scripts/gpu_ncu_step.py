"""Three warm-up training steps + one more, launched kernel by kernel (no graph) on the bench's model and a
Zipf-distributed batch -- the target of the ncu capture (skip the first 3 steps' launches)."""
import sys, torch
sys.path.insert(0, ".")
import bench
from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.runtime.engine import FusedEngine
B = 2048
cfg = DeepFMConfig(field_size=bench.FIELD, feature_size=bench.FEATURE, embedding_size=bench.EMB, deep_layers=bench.LAYERS,
                   dropout_keep=bench.KEEP, l2_reg=bench.L2)
eng = FusedEngine(cfg, B, optimizer="Adam", lr=bench.LR)
eng.load_from_module(DeepFM(cfg, seed=1))
pool, _ = bench.make_host_pool(B, 4, 0, 1)
eng.set_state(B, train=True)
torch.cuda.synchronize()
for i in range(4):
    slab, ids, vals, labels = pool[i]
    eng.train_step(ids.cuda(), vals.cuda(), labels.cuda(), set_state=False)
    torch.cuda.synchronize()
print("launches per step:", eng.launches_per_step())
