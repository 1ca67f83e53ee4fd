import copy, os, sys, tempfile
import torch
sys.path.insert(0, "/root/repo")
from deepfm_b200.data import tfrecord as T
from deepfm_b200.data.loader import BatchLoader
from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.runtime.trainer import Trainer, TrainerOptions
tmp = tempfile.mkdtemp(); path = os.path.join(tmp, "tr.tfrecords")
T.generate_synthetic(path, 256 * 40, 39, 117581, n_dense=13, seed=7)
B = 256
cfg = DeepFMConfig(39, 117581, 32, (256, 128, 64), (1.0,) * 3, l2_reg=0.0)
init = DeepFM(cfg, seed=1)
with torch.no_grad():
    init.fm_v.mul_(20); init.fm_w.mul_(20)
dev = torch.device("cuda", 0)
res = {}
for name, kw in (("graph", dict(backend="fused", cuda_graph=True)), ("plain", dict(backend="fused", cuda_graph=False)), ("eager", dict(backend="eager"))):
    tr = Trainer(cfg, TrainerOptions(batch_size=B, optimizer="GD", learning_rate=0.0, seed=1, **kw), device=dev, init_model=copy.deepcopy(init))
    loader = BatchLoader([path], 39, B, num_epochs=1, drop_remainder=True, num_workers=2, num_slots=6)
    losses = []
    for _ in tr.train_batches(loader):
        losses.append(tr.last_loss(wait_current=True))
    loader.close(); res[name] = torch.tensor(losses)
print("steps", len(res["graph"]))
print("graph-plain max", (res["graph"] - res["plain"]).abs().max().item())
print("graph-eager max", (res["graph"] - res["eager"]).abs().max().item())
print("plain-eager max", (res["plain"] - res["eager"]).abs().max().item())
print("graph", res["graph"][:8]); print("plain", res["plain"][:8]); print("eager", res["eager"][:8])
