"""BASELINE.json config 1: the reference script's defaults on ONE CPU worker.

    python scripts/config1_cpu.py [path/to/val.tfrecords] [--epochs 3] [--batch_size 256]

The reference's ``data/train.tfrecords`` is not in the tree (.MISSING_LARGE_BLOBS), so the bundled
``data/val.tfrecords`` (10,000 records) is split 80/20 in file order into ``tr.tfrecords`` / ``va.tfrecords`` and the
public CLI is run on it exactly as a user would (``python -m deepfm_b200.train --task_type train``, HVD.py:331-456):
script defaults -- 39 fields, 117,581 features, embedding 32, MLP 256-128-64, dropout keep 0.5, Adam lr 5e-4, l2 1e-4 --
per-epoch train -> evaluate, checkpoint, export.  Without the bundled file a synthetic Criteo-shaped one of the same
size stands in.  Prints one JSON line: examples/s over the whole train task (start-up, evaluation after every epoch
and export included), the pure training rate from the CLI's own log, and the final AUC.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("data", nargs="?", default="/root/reference/data/val.tfrecords")
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--batch_size", type=int, default=256)
    ap.add_argument("--threads", type=int, default=0, help="torch intra-op threads (0 = torch's default)")
    args = ap.parse_args()

    import deepfm_b200  # noqa: F401  (registers the package alias)
    from deepfm_b200.data import tfrecord as T

    work = tempfile.mkdtemp(prefix="dfm_config1_")
    src = args.data
    if not os.path.exists(src):
        src = os.path.join(work, "synthetic.tfrecords")
        T.generate_synthetic(src, 10000, 39, 117581, n_dense=13, seed=1, pos_rate=0.2553, zipf_s=1.05)
    recs = list(T.iter_records(src))
    cut = len(recs) * 8 // 10
    for name, part in (("tr.tfrecords", recs[:cut]), ("va.tfrecords", recs[cut:])):
        with T.TFRecordWriter(os.path.join(work, name)) as w:
            for r in part:
                w.write(r)

    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    if args.threads:
        env["OMP_NUM_THREADS"] = str(args.threads)
    cmd = [sys.executable, "-m", "deepfm_b200.train", "--task_type", "train", "--field_size", "39", "--feature_size",
           "117581", "--num_epochs", str(args.epochs), "--batch_size", str(args.batch_size), "--training_data_dir", work,
           "--val_data_dir", work, "--checkpoint_dir", os.path.join(work, "ckpt"), "--servable_model_dir",
           os.path.join(work, "export"), "--clear_existing_model", "True", "--log_steps", "10"]
    t0 = time.time()
    p = subprocess.run(cmd, env=env, capture_output=True, text=True)
    wall = time.time() - t0
    out = p.stdout + p.stderr
    if p.returncode != 0:
        print(out[-4000:])
        return p.returncode
    aucs = [float(x) for x in re.findall(r"'auc': ([0-9.]+)", out)]
    rates = [float(x) for x in re.findall(r"([0-9.]+) examples/sec", out)]
    steps = (cut // args.batch_size) * args.epochs
    print(json.dumps({"config": "script defaults, single CPU worker", "data": os.path.basename(src),
                      "train_records": cut, "val_records": len(recs) - cut, "epochs": args.epochs,
                      "batch_size": args.batch_size, "steps": steps,
                      "examples_per_s_whole_task": round(steps * args.batch_size / wall, 1), "wall_s": round(wall, 2),
                      "examples_per_s_training_only": rates, "auc_per_epoch": aucs,
                      "host_cores": os.cpu_count()}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
