"""AUC parity: the fused bf16 engine vs the eager fp32 restatement, same data, same initial weights.

    python scripts/auc_parity.py [path/to/val.tfrecords]     # default: synthetic Criteo-shaped data

Trains on the first 80% of the records for a few epochs and evaluates on the rest (tf.metrics.auc buckets).
On a box without CUDA only the eager arm runs."""
import copy, json, os, sys, tempfile, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepfm_b200.data import tfrecord as T
from deepfm_b200.data.loader import BatchLoader
from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.runtime.trainer import Trainer, TrainerOptions

src = sys.argv[1] if len(sys.argv) > 1 else None
tmp = tempfile.mkdtemp(prefix="dfm_auc_")
if src is None:
    src = os.path.join(tmp, "all.tfrecords")
    T.generate_synthetic(src, 50000, 39, 117581, n_dense=13, seed=7)
recs = list(T.iter_records(src))
cut = int(len(recs) * 0.8)
paths = {}
for name, part in (("tr", recs[:cut]), ("va", recs[cut:])):
    paths[name] = os.path.join(tmp, f"{name}.tfrecords")
    with T.TFRecordWriter(paths[name]) as w:
        for r in part:
            w.write(r)
B, EPOCHS = 256, int(os.environ.get("DFM_AUC_EPOCHS", "3"))
cfg = DeepFMConfig(39, 117581, 32, (256, 128, 64), (0.5, 0.5, 0.5), l2_reg=1e-4)
init = DeepFM(cfg, seed=1)
out = {"records": len(recs), "train": cut, "batch": B, "epochs": EPOCHS}
backends = ["eager"] + (["fused"] if torch.cuda.is_available() else [])
for backend in backends:
    dev = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")
    tr = Trainer(cfg, TrainerOptions(batch_size=B, optimizer="Adam", learning_rate=5e-4, backend=backend, seed=1),
                 device=dev, init_model=copy.deepcopy(init))   # the eager trainer trains the module it is given
    t0 = time.time()
    for _ in range(EPOCHS):
        loader = BatchLoader([paths["tr"]], 39, B, num_epochs=1, drop_remainder=True, num_workers=2, num_slots=6)
        for _hb in tr.train_batches(loader):
            pass
        loader.close()
    loader = BatchLoader([paths["va"]], 39, B, num_epochs=1, drop_remainder=False, num_workers=2)
    ev = tr.evaluate(loader, auc_mode="both")
    loader.close()
    out[backend] = {"auc": round(ev["auc"], 5), "auc_exact": round(ev.get("auc_exact", float("nan")), 5),
                    "loss": round(ev["data_loss"], 5), "train_s": round(time.time() - t0, 2), "device": str(dev)}
    tr.close()
if "fused" in out:
    out["auc_diff"] = round(abs(out["fused"]["auc"] - out["eager"]["auc"]), 5)
print(json.dumps(out))
