"""AUC two ways.

``tf.metrics.auc`` (reference HVD.py:271) is a streaming approximation: 200 thresholds
``{-1e-7, 1/199, ..., 198/199, 1+1e-7}``, per-threshold TP/FN/TN/FP counters, trapezoidal
area under (FPR, TPR).  ``StreamingAUC`` reproduces exactly that so numbers are comparable
with the reference's logs; ``exact_auc`` is the rank statistic (what sklearn reports).
The GPU path accumulates the same buckets with ``dfm_auc_hist`` (csrc/cuda/optim.cu).
"""
from __future__ import annotations

import numpy as np

K_EPS = 1e-7


def tf_thresholds(num_thresholds: int = 200) -> np.ndarray:
    t = [(i + 1) * 1.0 / (num_thresholds - 1) for i in range(num_thresholds - 2)]
    return np.array([0.0 - K_EPS] + t + [1.0 + K_EPS], dtype=np.float32)


def bucket_of(pred: np.ndarray, num_thresholds: int = 200) -> np.ndarray:
    """Number of thresholds strictly below each prediction (``pred > t`` in float32, as TF compares)."""
    th = tf_thresholds(num_thresholds)
    return (np.asarray(pred, dtype=np.float32)[:, None] > th[None, :]).sum(1)


def auc_from_hist(hist: np.ndarray) -> float:
    """hist[2, T+1]: bucket counts of positives (row 0) and negatives (row 1)."""
    hist = np.asarray(hist, dtype=np.float64)
    T = hist.shape[1] - 1
    # TP_j = #positives with pred > t_j = sum over buckets k > j
    tp = hist[0][::-1].cumsum()[::-1][1:]          # j = 0..T-1
    fp = hist[1][::-1].cumsum()[::-1][1:]
    fn = hist[0].sum() - tp
    tn = hist[1].sum() - fp
    eps = 1e-6
    tpr = (tp + eps) / (tp + fn + eps)
    fpr = fp / (fp + tn + eps)
    return float(((fpr[:T - 1] - fpr[1:]) * (tpr[:T - 1] + tpr[1:]) / 2.0).sum())


def tf_auc_numpy(labels: np.ndarray, preds: np.ndarray, num_thresholds: int = 200) -> float:
    labels = np.asarray(labels) > 0.5
    b = bucket_of(preds, num_thresholds)
    hist = np.zeros((2, num_thresholds + 1), dtype=np.int64)
    np.add.at(hist[0], b[labels], 1)
    np.add.at(hist[1], b[~labels], 1)
    return auc_from_hist(hist)


def exact_auc(labels: np.ndarray, preds: np.ndarray) -> float:
    """Mann-Whitney U with average ranks for ties."""
    labels = np.asarray(labels) > 0.5
    preds = np.asarray(preds, dtype=np.float64)
    n_pos, n_neg = int(labels.sum()), int((~labels).sum())
    if n_pos == 0 or n_neg == 0:
        return float("nan")
    order = np.argsort(preds, kind="mergesort")
    sp = preds[order]
    ranks = np.empty(len(preds), dtype=np.float64)
    i = 0
    n = len(sp)
    # average ranks over tie groups
    boundaries = np.flatnonzero(np.r_[True, sp[1:] != sp[:-1], True])
    for a, b in zip(boundaries[:-1], boundaries[1:]):
        ranks[order[a:b]] = 0.5 * (a + b - 1) + 1.0
    return float((ranks[labels].sum() - n_pos * (n_pos + 1) / 2.0) / (n_pos * n_neg))


def exact_auc_torch(labels, preds) -> float:
    """The same rank statistic computed where the tensors live (the GPU during fused evaluation): a device sort, tie
    groups by comparing neighbours, average ranks, float64 sums.  Nothing but the final scalar crosses to the host."""
    import torch
    preds = preds.reshape(-1).double()
    pos = labels.reshape(-1) > 0.5
    n = preds.numel()
    n_pos = int(pos.sum())
    n_neg = n - n_pos
    if n_pos == 0 or n_neg == 0:
        return float("nan")
    sp, order = torch.sort(preds, stable=True)
    new = torch.ones(n, dtype=torch.bool, device=preds.device)
    new[1:] = sp[1:] != sp[:-1]
    starts = torch.nonzero(new).reshape(-1)
    ends = torch.cat([starts[1:], torch.tensor([n], device=preds.device)])
    avg_rank = 0.5 * (starts + ends - 1).double() + 1.0
    ranks = avg_rank[torch.cumsum(new.long(), 0) - 1]
    s = ranks[pos[order]].sum()
    return float((s - n_pos * (n_pos + 1) / 2.0) / (float(n_pos) * float(n_neg)))


class StreamingAUC:
    """Accumulates over batches like the reference's eval_metric_ops (HVD.py:271-278)."""

    def __init__(self, num_thresholds: int = 200, keep_all: bool = True):
        self.T = num_thresholds
        self.hist = np.zeros((2, num_thresholds + 1), dtype=np.int64)
        self._all = ([], []) if keep_all else None

    def update(self, labels, preds) -> None:
        labels = np.asarray(labels, dtype=np.float32).reshape(-1)
        preds = np.asarray(preds, dtype=np.float32).reshape(-1)
        b = bucket_of(preds, self.T)
        pos = labels > 0.5
        np.add.at(self.hist[0], b[pos], 1)
        np.add.at(self.hist[1], b[~pos], 1)
        if self._all is not None:              # copies: callers pass views into ring buffers that are reused
            self._all[0].append(labels.copy())
            self._all[1].append(preds.copy())

    @property
    def keep_all_enabled(self) -> bool:
        return self._all is not None

    def remember(self, labels, preds) -> None:
        """Keeps raw (label, prediction) pairs for the exact AUC without touching the histogram."""
        if self._all is not None:
            self._all[0].append(np.asarray(labels, dtype=np.float32).reshape(-1).copy())
            self._all[1].append(np.asarray(preds, dtype=np.float32).reshape(-1).copy())

    def add_hist(self, hist) -> None:
        self.hist += np.asarray(hist, dtype=np.int64).reshape(2, self.T + 1)

    def result(self) -> float:
        return auc_from_hist(self.hist)

    def result_exact(self) -> float:
        if not self._all or not self._all[0]:
            return float("nan")
        return exact_auc(np.concatenate(self._all[0]), np.concatenate(self._all[1]))
