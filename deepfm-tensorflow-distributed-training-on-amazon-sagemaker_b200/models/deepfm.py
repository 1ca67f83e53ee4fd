"""DeepFM in plain PyTorch fp32 -- the numerical oracle and the CPU / fallback model.

An exact restatement of the reference ``model_fn`` (HVD.py:164-303, identical in PS.py:172-313):

    w      = fm_w[ids]                          # HVD.py:197
    y_w    = sum_f w * x                        # HVD.py:198
    e      = fm_v[ids] * x[..., None]           # HVD.py:201-203  (also the deep input)
    y_v    = 0.5 * sum_k[(sum_f e)^2 - sum_f e^2]   # HVD.py:204-206
    h_0    = reshape(e, [B, F*K])               # HVD.py:219-221
    h_l+1  = dropout_keep( BN( relu(h_l W_l + b_l) ) )   # FC -> ReLU -> BN -> dropout, HVD.py:223-235
    y_d    = h_L W_out + b_out                  # HVD.py:237-244
    y      = fm_bias + y_w + y_v + y_d          # HVD.py:247-249
    loss   = mean BCEWithLogits(y, label) + l2/2 |fm_w|^2 + l2/2 |fm_v|^2     # HVD.py:264-268

Variable names follow the TF graph so checkpoints interchange by name:
``fm_bias``, ``fm_w``, ``fm_v``, ``Deep-part/mlp{i}/{weights,biases}``,
``Deep-part/bn_{i}/{beta,gamma,moving_mean,moving_variance}``, ``Deep-part/deep_out/{weights,biases}``.
Quirks kept on purpose: ``--dropout`` values are keep probabilities (Q6), the MLP weights are
not regularised (Q2), BN sits after ReLU (Q5).
"""
from __future__ import annotations

import math
from collections import OrderedDict
from dataclasses import dataclass, field
from typing import Dict, List, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

BN_EPS = 1e-3  # tf.contrib.layers.batch_norm default epsilon


@dataclass
class DeepFMConfig:
    field_size: int
    feature_size: int
    embedding_size: int = 32
    deep_layers: Sequence[int] = (256, 128, 64)
    dropout_keep: Sequence[float] = (0.5, 0.5, 0.5)
    batch_norm: bool = False
    batch_norm_decay: float = 0.9
    l2_reg: float = 1e-4
    loss_type: str = "log_loss"

    def __post_init__(self):
        self.deep_layers = [int(x) for x in self.deep_layers]
        keep = [float(x) for x in self.dropout_keep]
        if len(keep) < len(self.deep_layers):
            keep = keep + [1.0] * (len(self.deep_layers) - len(keep))
        self.dropout_keep = keep[: len(self.deep_layers)]
        if self.loss_type not in ("log_loss", "square_loss"):
            raise ValueError(f"loss_type must be log_loss or square_loss, got {self.loss_type!r}")


def _glorot_normal_(t: torch.Tensor, fan_in: int, fan_out: int, gen=None) -> torch.Tensor:
    """tf.glorot_normal_initializer: truncated normal (+-2 sigma), variance 2/(fan_in+fan_out),
    with TF >= 1.14's 0.8796 correction for the truncation."""
    std = math.sqrt(2.0 / (fan_in + fan_out)) / 0.87962566103423978
    with torch.no_grad():
        nn.init.trunc_normal_(t, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=gen)
    return t


def _xavier_uniform_(t: torch.Tensor, fan_in: int, fan_out: int, gen=None) -> torch.Tensor:
    limit = math.sqrt(6.0 / (fan_in + fan_out))
    with torch.no_grad():
        t.uniform_(-limit, limit, generator=gen)
    return t


class DeepFM(nn.Module):
    def __init__(self, cfg: DeepFMConfig, seed: int | None = None):
        super().__init__()
        self.cfg = cfg
        gen = torch.Generator().manual_seed(seed) if seed is not None else None
        V, K, Fd = cfg.feature_size, cfg.embedding_size, cfg.field_size
        self.fm_bias = nn.Parameter(torch.zeros(1))
        self.fm_w = nn.Parameter(_glorot_normal_(torch.empty(V), V, V, gen))
        self.fm_v = nn.Parameter(_glorot_normal_(torch.empty(V, K), V, K, gen))
        self.mlp_w = nn.ParameterList()
        self.mlp_b = nn.ParameterList()
        self.bn_beta = nn.ParameterList()
        self.bn_gamma = nn.ParameterList()
        fan_in = Fd * K
        for i, n in enumerate(cfg.deep_layers):
            self.mlp_w.append(nn.Parameter(_xavier_uniform_(torch.empty(fan_in, n), fan_in, n, gen)))
            self.mlp_b.append(nn.Parameter(torch.zeros(n)))
            if cfg.batch_norm:
                self.bn_beta.append(nn.Parameter(torch.zeros(n)))
                self.bn_gamma.append(nn.Parameter(torch.ones(n)))
                self.register_buffer(f"bn_mean_{i}", torch.zeros(n))
                self.register_buffer(f"bn_var_{i}", torch.ones(n))
            fan_in = n
        self.out_w = nn.Parameter(_xavier_uniform_(torch.empty(fan_in, 1), fan_in, 1, gen))
        self.out_b = nn.Parameter(torch.zeros(1))

    # ------------------------------------------------------------------ forward
    def fm_terms(self, feat_ids: torch.Tensor, feat_vals: torch.Tensor):
        # embedding_lookup (HVD.py:197,201); F.embedding rather than advanced indexing: same values, a far cheaper
        # CPU gather and an index_add backward
        w = F.embedding(feat_ids, self.fm_w.unsqueeze(1)).squeeze(-1)
        y_w = (w * feat_vals).sum(1)
        e = F.embedding(feat_ids, self.fm_v) * feat_vals.unsqueeze(-1)
        y_v = 0.5 * (e.sum(1).pow(2) - e.pow(2).sum(1)).sum(1)
        return y_w, y_v, e

    def _bn(self, x: torch.Tensor, i: int, training: bool) -> torch.Tensor:
        mean_b, var_b = getattr(self, f"bn_mean_{i}"), getattr(self, f"bn_var_{i}")
        if training:
            mean = x.mean(0)
            var = x.var(0, unbiased=False)
            d = self.cfg.batch_norm_decay
            with torch.no_grad():   # in-place moving averages (updates_collections=None, HVD.py:312)
                mean_b.mul_(d).add_(mean.detach(), alpha=1 - d)
                var_b.mul_(d).add_(var.detach(), alpha=1 - d)
        else:
            mean, var = mean_b, var_b
        return (x - mean) * torch.rsqrt(var + BN_EPS) * self.bn_gamma[i] + self.bn_beta[i]

    # Optional: a callable (layer, rows, cols) -> bool keep mask that replaces torch's dropout stream in training --
    # used to give this oracle the fused engine's Philox masks (utils/philox.py) so the two differ by rounding only.
    dropout_mask_fn = None

    def forward(self, feat_ids: torch.Tensor, feat_vals: torch.Tensor, training: bool = False) -> torch.Tensor:
        """Returns logits [B]; ``prob = sigmoid(logits)`` (HVD.py:249)."""
        cfg = self.cfg
        feat_ids = feat_ids.long().view(-1, cfg.field_size)
        feat_vals = feat_vals.to(self._real()).view(-1, cfg.field_size)
        y_w, y_v, e = self.fm_terms(feat_ids, feat_vals)
        return self._tower(y_w, y_v, e, training)

    def forward_rows(self, w_rows: torch.Tensor, v_rows: torch.Tensor, inverse: torch.Tensor,
                     feat_vals: torch.Tensor, training: bool = False) -> torch.Tensor:
        """Same forward, but the table rows are passed in explicitly: ``w_rows[u]``, ``v_rows[u, K]`` are
        the rows of the batch's distinct ids and ``inverse[B, F]`` indexes into them.  Differentiating
        w.r.t. the rows yields the de-duplicated sparse gradient directly."""
        feat_vals = feat_vals.to(self._real()).view(-1, self.cfg.field_size)
        from ..ops import embed_fm_cpu
        if embed_fm_cpu.usable(w_rows, v_rows, inverse, feat_vals):
            # float32 on the host: one fused native pass each way instead of a dozen element-wise ops
            y_w, y_v, e = embed_fm_cpu.embed_fm(w_rows, v_rows, inverse, feat_vals)
            return self._tower(y_w, y_v, e, training)
        y_w = (F.embedding(inverse, w_rows.unsqueeze(1)).squeeze(-1) * feat_vals).sum(1)
        e = F.embedding(inverse, v_rows) * feat_vals.unsqueeze(-1)
        y_v = 0.5 * (e.sum(1).pow(2) - e.pow(2).sum(1)).sum(1)
        return self._tower(y_w, y_v, e, training)

    def _tower(self, y_w: torch.Tensor, y_v: torch.Tensor, e: torch.Tensor, training: bool) -> torch.Tensor:
        cfg = self.cfg
        h = e.reshape(e.shape[0], -1)
        for i in range(len(cfg.deep_layers)):
            h = F.relu(h @ self.mlp_w[i] + self.mlp_b[i])
            if cfg.batch_norm:
                h = self._bn(h, i, training)
            if training and cfg.dropout_keep[i] < 1.0:
                if self.dropout_mask_fn is not None:
                    h = h * self.dropout_mask_fn(i, h.shape[0], h.shape[1]).to(h.dtype) / cfg.dropout_keep[i]
                else:
                    h = F.dropout(h, p=1.0 - cfg.dropout_keep[i], training=True)
        y_d = (h @ self.out_w).view(-1) + self.out_b
        return self.fm_bias + y_w + y_v + y_d

    def _real(self) -> torch.dtype:
        """float32, or float64 when the module was converted with ``.double()`` (finite-difference checks)."""
        return torch.float64 if self.fm_v.dtype == torch.float64 else torch.float32

    def data_loss(self, logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        labels = labels.to(torch.float64 if logits.dtype == torch.float64 else torch.float32)
        if self.cfg.loss_type == "square_loss":
            return F.mse_loss(torch.sigmoid(logits), labels)
        return F.binary_cross_entropy_with_logits(logits, labels)

    def reg_loss(self) -> torch.Tensor:
        l2 = self.cfg.l2_reg
        return l2 * 0.5 * self.fm_w.pow(2).sum() + l2 * 0.5 * self.fm_v.pow(2).sum()

    def loss(self, logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        return self.data_loss(logits, labels) + self.reg_loss()

    # ------------------------------------------------------- TF variable naming
    def tf_variables(self) -> "OrderedDict[str, torch.Tensor]":
        """name -> tensor, names and shapes exactly as in the reference graph."""
        d: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        d["fm_bias"] = self.fm_bias
        d["fm_w"] = self.fm_w
        d["fm_v"] = self.fm_v
        for i in range(len(self.cfg.deep_layers)):
            d[f"Deep-part/mlp{i}/weights"] = self.mlp_w[i]
            d[f"Deep-part/mlp{i}/biases"] = self.mlp_b[i]
            if self.cfg.batch_norm:
                d[f"Deep-part/bn_{i}/beta"] = self.bn_beta[i]
                d[f"Deep-part/bn_{i}/gamma"] = self.bn_gamma[i]
                d[f"Deep-part/bn_{i}/moving_mean"] = getattr(self, f"bn_mean_{i}")
                d[f"Deep-part/bn_{i}/moving_variance"] = getattr(self, f"bn_var_{i}")
        d["Deep-part/deep_out/weights"] = self.out_w
        d["Deep-part/deep_out/biases"] = self.out_b
        return d

    def trainable_tf_variables(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((k, v) for k, v in self.tf_variables().items() if "moving_" not in k)

    def load_tf_variables(self, values: Dict[str, torch.Tensor], strict: bool = True) -> List[str]:
        """Copies by TF name; returns the names that were missing from ``values``."""
        missing = []
        with torch.no_grad():
            for name, dst in self.tf_variables().items():
                if name not in values:
                    missing.append(name)
                    continue
                src = torch.as_tensor(values[name]).to(dst.device, dst.dtype)
                if tuple(src.shape) != tuple(dst.shape):
                    raise ValueError(f"{name}: checkpoint shape {tuple(src.shape)} != model shape {tuple(dst.shape)}")
                dst.copy_(src)
        if strict and missing:
            raise KeyError(f"variables missing from checkpoint: {missing}")
        return missing

    def num_parameters(self) -> int:
        return sum(p.numel() for p in self.parameters())
