"""Debug/regression: Trainer.train_batches through the captured graph (pipelined prefetch) vs plain launches, same file,
same weights: parameters must agree up to atomic summation order after every epoch."""
import copy, os, sys, tempfile
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepfm_b200.data import tfrecord as T
from deepfm_b200.data.loader import BatchLoader
from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.runtime.trainer import Trainer, TrainerOptions

opt = sys.argv[1] if len(sys.argv) > 1 else "Adam"
keep = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
n_rec = int(sys.argv[3]) if len(sys.argv) > 3 else 40000
tmp = tempfile.mkdtemp(prefix="dfm_gvp_")
path = os.path.join(tmp, "tr.tfrecords")
T.generate_synthetic(path, n_rec, 39, 117581, n_dense=13, seed=7)
B = 256
cfg = DeepFMConfig(39, 117581, 32, (256, 128, 64), (keep,) * 3, l2_reg=1e-4)
init = DeepFM(cfg, seed=1)
dev = torch.device("cuda", 0)
mk = lambda graph: Trainer(cfg, TrainerOptions(batch_size=B, optimizer=opt, learning_rate=5e-4 if opt == "Adam" else 0.05,
                                               backend="fused", seed=1, cuda_graph=graph), device=dev, init_model=copy.deepcopy(init))
a, b = mk(True), mk(False)
c = Trainer(cfg, TrainerOptions(batch_size=B, optimizer=opt, learning_rate=5e-4 if opt == "Adam" else 0.05, backend="eager",
                                seed=1, dropout_stream="engine"), device=dev, init_model=copy.deepcopy(init))
for epoch in range(2):
    for tr in (a, b, c):
        loader = BatchLoader([path], 39, B, num_epochs=1, drop_remainder=True, num_workers=2, num_slots=6)
        n = sum(1 for _ in tr.train_batches(loader))
        loader.close()
    torch.cuda.synchronize()
    worst = 0.0
    for k, va in a.variables().items():
        vb = b.variables()[k]
        rel = ((va - vb).norm() / (vb.norm() + 1e-12)).item()
        worst = max(worst, rel)
        if rel > 1e-3:
            print(f"  epoch {epoch} {k}: rel diff {rel:.4g}")
    for name, x in (("graph", a), ("plain", b)):
        w = max(((x.variables()[k] - vc).norm() / (vc.norm() + 1e-12)).item() for k, vc in c.variables().items())
        print(f"  {name} vs eager fp32: worst rel diff {w:.3g}")
    sa, sb = a.engine.read_state(), b.engine.read_state()
    print(f"epoch {epoch}: {n} steps, worst rel diff {worst:.3g}; step counters {sa['step']} / {sb['step']}; lr {sa['lr']} / {sb['lr']}", flush=True)
