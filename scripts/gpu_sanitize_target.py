"""Small fused-engine workload for compute-sanitizer (memcheck / racecheck / synccheck / initcheck):
two training steps launched kernel by kernel, two graph replays with in-graph prefetch, one inference pass."""
import sys, torch
sys.path.insert(0, ".")
from deepfm_b200.data.loader import alloc_slab
from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.runtime.trainer import Trainer, TrainerOptions

B, F, V = 256, 39, 4000
cfg = DeepFMConfig(F, V, 32, (128, 64, 32), (0.5, 0.5, 0.5), l2_reg=1e-4)
g = torch.Generator().manual_seed(0)
slabs = []
for _ in range(3):
    slab, ids, vals, labels = alloc_slab(B, F, torch.int32, pin_memory=True)
    ids.copy_(torch.randint(1, V, (B, F), generator=g, dtype=torch.int32)); ids[:, :13] = torch.arange(1, 14, dtype=torch.int32)
    vals.copy_(torch.rand(B, F, generator=g)); labels.copy_((torch.rand(B, generator=g) < 0.3).float())
    slabs.append((slab, ids, vals, labels))
dev = torch.device("cuda", 0)
for graph in (False, True):
    tr = Trainer(cfg, TrainerOptions(batch_size=B, optimizer="Adam", learning_rate=1e-3, backend="fused", seed=1,
                                     cuda_graph=graph), device=dev, init_model=DeepFM(cfg, seed=1))
    tr.train_step_host(slabs[0][0], slabs[1][0] if graph else None)
    tr.train_step_host(slabs[1][0], None)
    loss = tr.last_loss(wait_current=True)
    prob = tr.engine.forward(slabs[2][1].cuda(), slabs[2][2].cuda(), slabs[2][3].cuda())
    torch.cuda.synchronize()
    print("graph" if graph else "eager", "loss %.5f" % loss, "prob mean %.4f" % prob[:B].mean().item())
    tr.close()
print("sanitize target done")
