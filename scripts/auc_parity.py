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
SEEDS = [int(x) for x in os.environ.get("DFM_AUC_SEEDS", "1,2,3").split(",")]
MODE = os.environ.get("DFM_AUC_TABLE_MODE", "lazy")
KEEP = float(os.environ.get("DFM_AUC_KEEP", "0.5"))
OPT = os.environ.get("DFM_AUC_OPT", "Adam")
LR = float(os.environ.get("DFM_AUC_LR", "5e-4"))
GRAPH = os.environ.get("DFM_AUC_GRAPH", "1") != "0"
cfg = DeepFMConfig(39, 117581, 32, (256, 128, 64), (KEEP, KEEP, KEEP), l2_reg=float(os.environ.get("DFM_AUC_L2", "1e-4")))
out = {"keep": KEEP, "opt": OPT, "lr": LR, "graph": GRAPH,"records": len(recs), "train": cut, "batch": B, "epochs": EPOCHS, "seeds": SEEDS, "table_update_mode": MODE,
       "dropout": "the eager arm uses the fused engine's Philox keep masks (same seed, step, layer, row, column)", "runs": []}
backends = ["eager"] + (["fused"] if torch.cuda.is_available() else [])
for seed in SEEDS:
    init = DeepFM(cfg, seed=seed)
    run = {"seed": seed}
    for backend in backends:
        dev = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")
        tr = Trainer(cfg, TrainerOptions(batch_size=B, optimizer=OPT, learning_rate=LR, backend=backend, seed=seed,
                                         table_update_mode=MODE, dropout_stream="engine", cuda_graph=GRAPH),
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
        run[backend] = {"auc": round(ev["auc"], 5), "auc_exact": round(ev.get("auc_exact", float("nan")), 5),
                        "loss": round(ev["data_loss"], 5), "train_s": round(time.time() - t0, 2)}
        tr.close()
    if "fused" in run:
        run["auc_diff"] = round(run["fused"]["auc"] - run["eager"]["auc"], 5)
    out["runs"].append(run)
if "fused" in backends:
    d = [r["auc_diff"] for r in out["runs"]]
    out["auc_diff_mean"] = round(sum(d) / len(d), 5)
    out["auc_diff_max_abs"] = round(max(abs(x) for x in d), 5)
for b in backends:
    a = [r[b]["auc"] for r in out["runs"]]
    out[f"{b}_auc_mean"] = round(sum(a) / len(a), 5)
    out[f"{b}_auc_spread"] = round(max(a) - min(a), 5)
print(json.dumps(out))
