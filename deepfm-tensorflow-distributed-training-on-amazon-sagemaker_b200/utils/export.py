"""Servable export and its loader -- the SavedModel step of the reference.

``DeepFM.export_savedmodel(servable_model_dir, raw_serving_input_receiver_fn)`` (HVD.py:477-493,
PS.py:535-551) writes ``<dir>/<unix_ts>/saved_model.pb + variables/`` with one signature,
``serving_default``: inputs ``feat_ids`` int64 ``[None, F]`` and ``feat_vals`` float32 ``[None, F]``,
output ``prob`` float32 ``[None]`` (HVD.py:252-256, 478-483).

The same contract here, as a self-contained directory that needs nothing but this package (or
just numpy, via the documented layout) to serve:

    <servable_model_dir>/<unix_ts>/
        signature.json           inputs/outputs of ``serving_default`` + model hyper-parameters
        variables/variables.index, variables.data-00000-of-00001   TF V2 bundle, reference names

``load_servable(dir)`` returns a callable with that signature.
"""
from __future__ import annotations

import json
import os
import time
from typing import Callable, Dict, Optional

import numpy as np
import torch

from ..models.deepfm import DeepFM, DeepFMConfig
from .checkpoint import read_bundle, write_bundle

SIGNATURE_KEY = "serving_default"


def export_servable(model: DeepFM, servable_model_dir: str, timestamp: Optional[int] = None) -> str:
    ts = int(time.time()) if timestamp is None else int(timestamp)
    out = os.path.join(servable_model_dir, str(ts))
    while os.path.exists(out):       # export_savedmodel never overwrites an export
        ts += 1
        out = os.path.join(servable_model_dir, str(ts))
    os.makedirs(os.path.join(out, "variables"))
    cfg = model.cfg
    tensors = {k: v.detach().cpu().numpy() for k, v in model.tf_variables().items()}
    write_bundle(os.path.join(out, "variables", "variables"), tensors)
    sig = {
        "signature_def": {SIGNATURE_KEY: {
            "method_name": "tensorflow/serving/predict",
            "inputs": {"feat_ids": {"dtype": "int64", "shape": [None, cfg.field_size]},
                       "feat_vals": {"dtype": "float32", "shape": [None, cfg.field_size]}},
            "outputs": {"prob": {"dtype": "float32", "shape": [None]}}}},
        "model": {"field_size": cfg.field_size, "feature_size": cfg.feature_size,
                  "embedding_size": cfg.embedding_size, "deep_layers": list(cfg.deep_layers),
                  "dropout_keep": list(cfg.dropout_keep), "batch_norm": cfg.batch_norm,
                  "batch_norm_decay": cfg.batch_norm_decay, "l2_reg": cfg.l2_reg, "loss_type": cfg.loss_type},
        "framework": "deepfm_b200", "format_version": 1,
    }
    with open(os.path.join(out, "signature.json"), "w") as f:
        json.dump(sig, f, indent=1)
    return out


def latest_export(servable_model_dir: str) -> Optional[str]:
    if not os.path.isdir(servable_model_dir):
        return None
    stamps = [d for d in os.listdir(servable_model_dir) if d.isdigit()]
    return os.path.join(servable_model_dir, max(stamps, key=int)) if stamps else None


def load_servable(path: str, device: str = "cpu", backend: str = "eager",
                  max_batch: int = 4096) -> Callable[[np.ndarray, np.ndarray], Dict[str, np.ndarray]]:
    """``path`` is an export directory or its parent (then the newest export is used).

    ``backend="fused"`` (CUDA device): requests are answered by the sm_100a inference kernels -- fused embedding gather +
    FM, tcgen05 MLP layers -- in chunks of at most ``max_batch`` rows, instead of eager PyTorch ops."""
    if not os.path.exists(os.path.join(path, "signature.json")):
        newest = latest_export(path)
        if newest is None:
            raise FileNotFoundError(f"no servable export under {path}")
        path = newest
    sig = json.load(open(os.path.join(path, "signature.json")))
    m = sig["model"]
    cfg = DeepFMConfig(m["field_size"], m["feature_size"], m["embedding_size"], m["deep_layers"], m["dropout_keep"],
                       m["batch_norm"], m["batch_norm_decay"], m["l2_reg"], m["loss_type"])
    model = DeepFM(cfg, seed=0)
    tensors = read_bundle(os.path.join(path, "variables", "variables"))
    model.load_tf_variables({k: torch.from_numpy(v) for k, v in tensors.items()})
    if backend == "fused":
        from ..runtime.engine import FusedEngine
        dev = torch.device(device if str(device).startswith("cuda") else "cuda")
        engine = FusedEngine(cfg, max_batch, optimizer="GD", device=dev)
        engine.load_from_module(model)

        def serve_fused(feat_ids, feat_vals) -> Dict[str, np.ndarray]:
            raw = np.asarray(feat_ids)
            if raw.size and (int(raw.min()) < 0 or int(raw.max()) >= cfg.feature_size):     # before narrowing to int32
                raise IndexError(f"feat_ids must lie in [0, feature_size = {cfg.feature_size}) "
                                 "(tf.nn.embedding_lookup raises InvalidArgument there)")
            ids = torch.as_tensor(np.ascontiguousarray(raw, dtype=np.int32)).reshape(-1, cfg.field_size)
            vals = torch.as_tensor(np.ascontiguousarray(feat_vals, dtype=np.float32)).reshape(-1, cfg.field_size)
            out = torch.empty(ids.shape[0], dtype=torch.float32)
            for o in range(0, ids.shape[0], max_batch):
                i = ids[o:o + max_batch].to(dev).contiguous()
                v = vals[o:o + max_batch].to(dev).contiguous()
                out[o:o + i.shape[0]] = engine.forward(i, v).float().cpu()
            return {"prob": out.numpy()}

        serve_fused.config = cfg
        serve_fused.model = model
        serve_fused.engine = engine
        return serve_fused
    if backend != "eager":
        raise ValueError(f"load_servable: unknown backend {backend!r} (eager | fused)")
    model = model.to(device).eval()

    def serve(feat_ids, feat_vals) -> Dict[str, np.ndarray]:
        with torch.no_grad():
            ids = torch.as_tensor(np.asarray(feat_ids, dtype=np.int64), device=device)
            vals = torch.as_tensor(np.asarray(feat_vals, dtype=np.float32), device=device)
            if ids.numel() and (int(ids.min()) < 0 or int(ids.max()) >= cfg.feature_size):
                raise IndexError(f"feat_ids must lie in [0, feature_size = {cfg.feature_size}) "
                                 "(tf.nn.embedding_lookup raises InvalidArgument there)")
            if ids.device.type == "cpu" and ids.dim() == 2:
                # the whole table as "rows", the ids as the index: float32 on the host takes the fused native gather + FM
                logits = model.forward_rows(model.fm_w, model.fm_v, ids, vals, training=False)
            else:
                logits = model(ids, vals, training=False)
            return {"prob": torch.sigmoid(logits).cpu().numpy()}

    serve.config = cfg
    serve.model = model
    return serve
