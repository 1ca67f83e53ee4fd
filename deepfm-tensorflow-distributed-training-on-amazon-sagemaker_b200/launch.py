"""Local stand-in for the SageMaker estimators of the two notebooks.

``TensorFlow(entry_point=..., distribution={'mpi': {'processes_per_host': 4}}, hyperparameters=...,
input_mode='File'|'Pipe').fit({'training': s3://..., 'evaluation': s3://...})`` (NB-HVD:109-155) and the
``{'parameter_server': {'enabled': True}}`` variant (NB-PS:97-143) become, on one box:

    python -m deepfm_b200.launch --distribution mpi --processes_per_host 8 \\
        --hyperparameters feature_size=117581 field_size=39 batch_size=1024 ... \\
        --channels training=/data/train evaluation=/data/val [--input_mode Pipe]

What SageMaker's container does for a training script is reproduced here: the hyper-parameters arrive as
``--key value`` arguments, the channels as ``SM_CHANNELS`` / ``SM_CHANNEL_<NAME>`` (File mode: directories,
Pipe mode: FIFOs ``<root>/<channel>_<epoch>`` fed with the channel's TFRecord files once per epoch),
``SM_HOSTS`` / ``SM_CURRENT_HOST`` name the cluster, and the processes are started one per GPU with the
rendezvous variables torchrun would set (``mpi``) or with one PS + one worker task per "host" and a
``TF_CONFIG`` each (``parameter_server``)."""
from __future__ import annotations

import argparse
import glob
import json
import os
import subprocess
import sys
import tempfile
import threading
from typing import Dict, List


def _parse_kv(items: List[str]) -> Dict[str, str]:
    out = {}
    for it in items or []:
        if "=" not in it:
            raise SystemExit(f"expected key=value, got {it!r}")
        k, v = it.split("=", 1)
        out[k] = v
    return out


def _feed_fifo(path: str, files: List[str]) -> None:
    """One epoch of a pipe-mode channel: the channel's files, concatenated, through a FIFO."""
    with open(path, "wb") as fo:
        for f in files:
            with open(f, "rb") as fi:
                while True:
                    chunk = fi.read(1 << 20)
                    if not chunk:
                        break
                    fo.write(chunk)


def _pipe_channels(channels: Dict[str, str], epochs: int, root: str) -> List[threading.Thread]:
    threads = []
    os.makedirs(root, exist_ok=True)
    for name, src in channels.items():
        files = sorted(glob.glob(os.path.join(src, "**", "*.tfrecords"), recursive=True)) if os.path.isdir(src) else [src]
        for e in range(epochs):
            fifo = os.path.join(root, f"{name}_{e}")
            if not os.path.exists(fifo):
                os.mkfifo(fifo)
            t = threading.Thread(target=_feed_fifo, args=(fifo, files), daemon=True)
            t.start()
            threads.append(t)
    return threads


def build_commands(args) -> List[dict]:
    """One dict per process: {'argv': [...], 'env': {...}} (split out so tests can inspect it)."""
    hp = _parse_kv(args.hyperparameters)
    channels = _parse_kv(args.channels)
    hosts = [f"algo-{i + 1}" for i in range(args.instance_count)]
    base_env = {
        "SM_HOSTS": json.dumps(hosts),
        "SM_CHANNELS": json.dumps(sorted(channels)),         # SageMaker sorts them: 'evaluation' first (README.md:81)
        "SM_NUM_GPUS": str(args.processes_per_host),
        "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(args.master_port),
    }
    pipe = args.input_mode.lower() == "pipe"
    if pipe:
        base_env["SM_INPUT_DATA_DIR"] = args.pipe_root
        hp.setdefault("pipe_mode", "1")
    else:
        for name, path in channels.items():
            base_env[f"SM_CHANNEL_{name.upper()}"] = path
        if "training" in channels:
            hp.setdefault("training_data_dir", channels["training"])
        if "evaluation" in channels:
            hp.setdefault("val_data_dir", channels["evaluation"])
    argv_hp = [x for k, v in hp.items() for x in (f"--{k}", str(v))]
    entry = [sys.executable, "-m", "deepfm_b200"]
    procs = []
    if args.distribution == "mpi":
        world = args.instance_count * args.processes_per_host
        for h, host in enumerate(hosts):
            for lr in range(args.processes_per_host):
                rank = h * args.processes_per_host + lr
                env = dict(base_env, SM_CURRENT_HOST=host, RANK=str(rank), LOCAL_RANK=str(lr), WORLD_SIZE=str(world),
                           LOCAL_WORLD_SIZE=str(args.processes_per_host))
                procs.append({"argv": entry + ["--dist", "hvd", "--worker_per_host", str(args.processes_per_host)] + argv_hp,
                              "env": env})
    else:
        # SageMaker's parameter_server distribution: every instance runs one PS task and one worker task
        ps_hosts = ",".join(f"{h}:2223" for h in hosts)
        worker_hosts = ",".join(f"{h}:2222" for h in hosts)
        from .parallel.env import set_dist_env
        for job, n in (("ps", len(hosts)), ("worker", len(hosts))):
            for i in range(n):
                saved = os.environ.pop("TF_CONFIG", None)
                tf_config = set_dist_env(1, ps_hosts, worker_hosts, job, i)
                os.environ.pop("TF_CONFIG", None)
                if saved is not None:
                    os.environ["TF_CONFIG"] = saved
                env = dict(base_env, SM_CURRENT_HOST=hosts[i], TF_CONFIG=json.dumps(tf_config))
                procs.append({"argv": entry + ["--dist", "ps", "--dist_mode", "1"] + argv_hp, "env": env})
    return procs


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--distribution", choices=["mpi", "parameter_server"], default="mpi")
    ap.add_argument("--instance_count", type=int, default=1, help="simulated hosts (all on this box)")
    ap.add_argument("--processes_per_host", type=int, default=1, help="mpi: ranks (GPUs) per host, NB-HVD:91")
    ap.add_argument("--input_mode", choices=["File", "Pipe", "file", "pipe"], default="File")
    ap.add_argument("--hyperparameters", nargs="*", default=[], metavar="key=value")
    ap.add_argument("--channels", nargs="*", default=[], metavar="name=path")
    ap.add_argument("--pipe_root", default="", help="directory for the pipe-mode FIFOs (default: a temp dir)")
    ap.add_argument("--master_port", type=int, default=29555)
    ap.add_argument("--dry_run", action="store_true", help="print the processes instead of starting them")
    args = ap.parse_args(argv)
    tmp = None
    if args.input_mode.lower() == "pipe" and not args.pipe_root:
        tmp = tempfile.TemporaryDirectory(prefix="dfm_pipe_")
        args.pipe_root = tmp.name
    procs = build_commands(args)
    if args.dry_run:
        for p in procs:
            print(json.dumps(p))
        return 0
    if args.input_mode.lower() == "pipe":
        epochs = int(_parse_kv(args.hyperparameters).get("num_epochs", 1))
        _pipe_channels(_parse_kv(args.channels), epochs, args.pipe_root)
    running = [subprocess.Popen(p["argv"], env=dict(os.environ, **p["env"])) for p in procs]
    rc = 0
    for r in running:
        rc = r.wait() or rc
    if tmp is not None:
        tmp.cleanup()
    return rc


if __name__ == "__main__":
    sys.exit(main())
