"""ncu target for the two collective kernels: 8 virtual ranks on one GPU (runtime.engine.EmulatedRanks), the bench's
model and batch size per rank, four steps.  Under ncu every launch is replayed in isolation, which a real multi-GPU
rank could not survive (it waits for its peers' flags); here all ranks are inside the one cooperative launch."""
import sys, torch
sys.path.insert(0, ".")
import bench
from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.runtime.engine import EmulatedRanks
N, B = 8, 2048
cfg = DeepFMConfig(field_size=bench.FIELD, feature_size=bench.FEATURE, embedding_size=bench.EMB, deep_layers=bench.LAYERS,
                   dropout_keep=bench.KEEP, l2_reg=bench.L2)
emu = EmulatedRanks(cfg, N, B, optimizer="Adam", lr=bench.LR)
emu.load_from_module(DeepFM(cfg, seed=1))
pool, _ = bench.make_host_pool(B, N * 4, 0, 1)
torch.cuda.synchronize()
for step in range(4):
    batches = [tuple(t.cuda() for t in pool[step * N + r][1:]) for r in range(N)]
    emu.train_step(batches)
    torch.cuda.synchronize()
print("emulated steps done; ctas per rank:", 148 // N)
