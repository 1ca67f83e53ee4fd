"""Command-line surface of the reference scripts, without absl/TensorFlow.

Every flag of ``2-hvd-gpu/DeepFM-hvd-tfrecord-vectorized-map.py`` (HVD.py:36-98, 28 flags) and
``1-ps-cpu/DeepFM-dist-ps-for-multipleCPU-multiInstance.py`` (PS.py:37-107, 34 flags) is
accepted with the same name, type and default, so a SageMaker ``hyperparameters``
dict or an existing launch line keeps working.  Differences, all deliberate:

* environment is optional: ``SM_HOSTS`` / ``SM_CURRENT_HOST`` / ``SM_CHANNELS`` fall back to
  torchrun variables or localhost instead of crashing at import time (HVD.py:71);
* unknown ``--key value`` pairs are ignored with a note, as ``tf.app.run`` did (the notebooks
  pass ``perform_shuffle``, NB-HVD:104);
* booleans accept ``--flag``, ``--noflag``, ``--flag=true|false`` and SageMaker's two-token
  ``--flag True`` form;
* ``GD`` works and ``optimizer`` is case-insensitive (the reference crashes on ``GD``);
* new flags (``--table_update_mode``, ``--backend``, ``--drop_remainder_eval`` ...) are listed at
  the end of FLAG_SPECS.
"""
from __future__ import annotations

import json
import os
import sys
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple


def _env_hosts() -> List[str]:
    raw = os.environ.get("SM_HOSTS")
    if raw:
        try:
            v = json.loads(raw)
            if isinstance(v, list) and v:
                return [str(x) for x in v]
        except ValueError:
            return [h for h in raw.split(",") if h]
    return ["localhost"]


def _env_current_host() -> str:
    return os.environ.get("SM_CURRENT_HOST") or _env_hosts()[0]


def _env_int(name: str, default: int) -> int:
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


# name -> (type, default, help).  type in {int, float, str, bool, "list"}
FLAG_SPECS: Dict[str, Tuple[Any, Any, str]] = {
    # ---- model / optimisation (both scripts) ----
    "feature_size": (int, 0, "Number of features (rows of fm_w / fm_v)"),
    "field_size": (int, 0, "Number of fields"),
    "embedding_size": (int, 32, "Embedding size"),
    "num_epochs": (int, 10, "Number of epochs"),
    "batch_size": (int, 64, "Batch size per worker"),
    "log_steps": (int, 1000, "log every N steps (inert in the reference; honoured here)"),
    "learning_rate": (float, 0.0005, "learning rate (multiplied by world size in sync-DP mode, HVD.py:171)"),
    "l2_reg": (float, 0.0001, "L2 regularization on fm_w and fm_v"),
    "loss_type": (str, "log_loss", "loss type {square_loss, log_loss}"),
    "optimizer": (str, "Adam", "optimizer type {Adam, Adagrad, GD, Momentum, ftrl}"),
    "deep_layers": (str, "256,128,64", "deep layers"),
    "dropout": (str, "0.5,0.5,0.5", "dropout KEEP probabilities per deep layer (HVD.py:235)"),
    "batch_norm": (bool, False, "perform batch normalization"),
    "batch_norm_decay": (float, 0.9, "decay for the moving average"),
    # ---- data / output locations ----
    "training_data_dir": (str, "", "training data dir (globbed for **/tr*.tfrecords)"),
    "val_data_dir": (str, "", "validation data dir (**/va*.tfrecords, **/te*.tfrecords; pred.txt goes here)"),
    "checkpoint_dir": (str, "", "local checkpoint dir (HVD script)"),
    "model_dir": (str, "", "model/checkpoint dir (PS script; alias of checkpoint_dir)"),
    "servable_model_dir": (str, "", "export servable model here"),
    "task_type": (str, "train", "task type {train, infer, eval, export}"),
    "clear_existing_model": (bool, False, "clear existing model or not"),
    # ---- cluster description ----
    "hosts": ("list", None, "all cluster instance names (default: $SM_HOSTS or localhost)"),
    "current_host": (str, None, "this host's name (default: $SM_CURRENT_HOST)"),
    "pipe_mode": (int, 0, "1 = stream records from FIFOs (SageMaker pipe mode)"),
    "worker_per_host": (int, 1, "worker processes per training instance"),
    "training_channel_name": (str, "", "training channel name"),
    "evaluation_channel_name": (str, "", "evaluation channel name"),
    "enable_s3_shard": (bool, False, "data already sharded per host (ShardedByS3Key)"),
    "enable_data_multi_path": (bool, False, "a different dataset path per channel (pipe mode)"),
    # ---- PS script extras ----
    "dist_mode": (int, 0, "distribution mode {0-single, 1-single dist, 2-multi dist}"),
    "ps_hosts": (str, "", "comma-separated parameter-server hosts"),
    "worker_hosts": (str, "", "comma-separated worker hosts"),
    "job_name": (str, "", "job name: worker, ps or master"),
    "task_index": (int, 0, "index of task within the job"),
    "num_threads": (int, 16, "number of host threads"),
    "num_GPUs": (int, None, "GPUs on this host (default: $SM_NUM_GPUS)"),
    "num_CPUs": (int, None, "CPUs on this host (default: $SM_NUM_CPUS)"),
    # ---- additions of this framework ----
    "dist": (str, "auto", "distribution back-end {auto, none, hvd, ps}: hvd = synchronous data parallel"),
    "backend": (str, "auto", "compute path {auto, fused, eager}: fused = sm_100a kernels"),
    "table_update_mode": (str, "lazy", "{lazy, dense_exact}: embedding-table update semantics (SURVEY Q3)"),
    "drop_remainder_eval": (bool, False, "drop the tail batch in eval/infer like the reference (SURVEY Q4)"),
    "auc_mode": (str, "tf", "{tf, exact, both}: 200-threshold tf.metrics.auc or exact rank AUC"),
    "seed": (int, 0, "RNG seed (init, dropout, file shuffle)"),
    "loader_threads": (int, 0, "TFRecord parser threads (0: the cores this process may use / ranks on this host, 2..16)"),
    "prefetch_batches": (int, 8, "ring depth of pinned batch buffers"),
    "save_checkpoints_secs": (int, 600, "checkpoint period in seconds (Estimator default)"),
    "keep_checkpoint_max": (int, 5, "checkpoints kept (Estimator default)"),
    "max_steps": (int, 0, "stop after this many optimizer steps (0 = run all epochs)"),
    "cuda_graph": (bool, True, "replay the fused step as a CUDA graph"),
    "ps_eval_throttle_secs": (int, 1200, "PS mode: minimum seconds between evaluations (PS.py:516)"),
    "ps_eval_start_delay_secs": (int, 1000, "PS mode: delay before the first evaluation (PS.py:515)"),
}

_TRUE = {"true", "1", "yes", "y", "t", "on"}
_FALSE = {"false", "0", "no", "n", "f", "off"}


class Flags:
    """Attribute bag with the reference's flag names."""

    def __init__(self, values: Dict[str, Any]):
        self.__dict__.update(values)

    def as_dict(self) -> Dict[str, Any]:
        return dict(self.__dict__)

    @property
    def deep_layer_list(self) -> List[int]:
        return [int(x) for x in str(self.deep_layers).split(",") if x.strip()]

    @property
    def dropout_keep_list(self) -> List[float]:
        return [float(x) for x in str(self.dropout).split(",") if x.strip()]

    @property
    def parser_threads(self) -> int:
        """--loader_threads, or a share of the cores this process may run on (tf.data AUTOTUNE's role, HVD.py:159)."""
        if int(self.loader_threads) > 0:
            return int(self.loader_threads)
        cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 4)
        local = os.environ.get("LOCAL_WORLD_SIZE") or os.environ.get("OMPI_COMM_WORLD_LOCAL_SIZE") or "1"
        return max(2, min(16, cores // max(1, int(local))))

    @property
    def effective_model_dir(self) -> str:
        return self.checkpoint_dir or self.model_dir

    def __repr__(self) -> str:
        return "Flags(" + ", ".join(f"{k}={v!r}" for k, v in sorted(self.__dict__.items())) + ")"


def _convert(name: str, typ: Any, raw: str) -> Any:
    try:
        if typ is int:
            return int(float(raw)) if raw.strip() else 0
        if typ is float:
            return float(raw)
        if typ == "list":
            raw = raw.strip()
            if raw.startswith("["):
                return [str(x) for x in json.loads(raw)]
            return [x for x in raw.split(",") if x]
        return raw
    except ValueError as e:
        raise SystemExit(f"--{name}: cannot parse {raw!r} ({e})")


def parse_flags(argv: Optional[Sequence[str]] = None) -> Tuple[Flags, List[str]]:
    """Returns (flags, ignored_tokens)."""
    argv = list(sys.argv[1:] if argv is None else argv)
    values: Dict[str, Any] = {k: spec[1] for k, spec in FLAG_SPECS.items()}
    ignored: List[str] = []
    i = 0
    while i < len(argv):
        tok = argv[i]
        i += 1
        if not tok.startswith("-") or tok in ("-", "--"):
            ignored.append(tok)
            continue
        body = tok.lstrip("-")
        name, eq, val = body.partition("=")
        if name in FLAG_SPECS:
            typ = FLAG_SPECS[name][0]
            if typ is bool:
                if eq:
                    low = val.lower()
                elif i < len(argv) and argv[i].lower() in _TRUE | _FALSE:
                    low = argv[i].lower()   # SageMaker: "--flag True"
                    i += 1
                else:
                    low = "true"
                if low not in _TRUE | _FALSE:
                    raise SystemExit(f"--{name}: expected a boolean, got {val!r}")
                values[name] = low in _TRUE
            else:
                if not eq:
                    if i >= len(argv):
                        raise SystemExit(f"--{name}: missing value")
                    val = argv[i]
                    i += 1
                values[name] = _convert(name, typ, val)
        elif name.startswith("no") and name[2:] in FLAG_SPECS and FLAG_SPECS[name[2:]][0] is bool and not eq:
            values[name[2:]] = False
        else:
            # unknown flag: swallow its value token too when it looks like "--key value"
            ignored.append(tok)
            if not eq and i < len(argv) and not argv[i].startswith("--"):
                ignored.append(argv[i])
                i += 1
    if values["hosts"] is None:
        values["hosts"] = _env_hosts()
    if values["current_host"] is None:
        values["current_host"] = _env_current_host()
    if values["num_GPUs"] is None:
        values["num_GPUs"] = _env_int("SM_NUM_GPUS", 0)
    if values["num_CPUs"] is None:
        values["num_CPUs"] = _env_int("SM_NUM_CPUS", os.cpu_count() or 1)
    f = Flags(values)
    validate_flags(f)
    return f, ignored


def validate_flags(f: Flags) -> None:
    from .optimizers import canonical_optimizer
    canonical_optimizer(f.optimizer)
    if f.task_type not in ("train", "eval", "infer", "export"):
        raise SystemExit(f"--task_type must be one of train, eval, infer, export (got {f.task_type!r})")
    if f.loss_type not in ("log_loss", "square_loss"):
        raise SystemExit(f"--loss_type must be log_loss or square_loss (got {f.loss_type!r})")
    if f.table_update_mode not in ("lazy", "dense_exact"):
        raise SystemExit("--table_update_mode must be lazy or dense_exact")
    if f.backend not in ("auto", "fused", "eager"):
        raise SystemExit("--backend must be auto, fused or eager")
    if f.dist not in ("auto", "none", "hvd", "ps"):
        raise SystemExit("--dist must be auto, none, hvd or ps")
    if f.field_size <= 0 or f.feature_size <= 0:
        raise SystemExit("--field_size and --feature_size are required (e.g. 39 and 117581)")
    if len(f.dropout_keep_list) < len(f.deep_layer_list):
        raise SystemExit("--dropout needs one keep probability per deep layer")


def channel_names() -> List[str]:
    """$SM_CHANNELS (alphabetically sorted by SageMaker: 'evaluation' first, README.md:81)."""
    raw = os.environ.get("SM_CHANNELS")
    if not raw:
        return []
    try:
        return [str(x) for x in json.loads(raw)]
    except ValueError:
        return [x for x in raw.split(",") if x]


def print_flags(f: Flags, out=None) -> None:
    """The reference's start-up dump (HVD.py:334-350)."""
    out = out or sys.stdout
    for k in ("task_type", "checkpoint_dir", "training_data_dir", "val_data_dir", "num_epochs", "feature_size",
              "field_size", "embedding_size", "batch_size", "deep_layers", "dropout", "loss_type", "optimizer",
              "learning_rate", "batch_norm_decay", "batch_norm", "l2_reg"):
        print(f"{k} ", getattr(f, k), file=out)
