"""The five update rules of the reference, with TF1's exact formulas and slot names.

    Adam      tf.train.AdamOptimizer(lr, beta1=0.9, beta2=0.999, epsilon=1e-8)        HVD.py:282-284
    Adagrad   tf.train.AdagradOptimizer(lr, initial_accumulator_value=1e-8)          HVD.py:286-288
    Momentum  tf.train.MomentumOptimizer(lr, momentum=0.95)                          HVD.py:290-292
    ftrl      tf.train.FtrlOptimizer(lr)  (lr_power -0.5, accum 0.1, l1=l2=0)        HVD.py:294
    GD        tf.train.GradientDescentOptimizer(lr) -- named in the reference's help text but
              left unimplemented there; implemented here.

This is the eager / CPU implementation (oracle for the CUDA kernels in csrc/cuda/optim.cu, the
optimizer of the parameter-server mode).  Works on whole tensors (``apply_dense``) or on a set
of distinct rows (``apply_rows``: the lazy sparse update of the embedding tables).
Slots are kept under their TF checkpoint names (``<var>/Adam``, ``<var>/Adam_1``, ...).
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch

OPT_IDS = {"adam": 0, "adagrad": 1, "momentum": 2, "ftrl": 3, "gd": 4}
_SLOT_NAMES = {
    "adam": ("Adam", "Adam_1"),
    "adagrad": ("Adagrad", None),
    "momentum": ("Momentum", None),
    "ftrl": ("Ftrl", "Ftrl_1"),
    "gd": (None, None),
}
_SLOT_INIT = {"adam": (0.0, 0.0), "adagrad": (1e-8, 0.0), "momentum": (0.0, 0.0), "ftrl": (0.1, 0.0), "gd": (0.0, 0.0)}


_cpu_lib = None


def _native_adam():
    """libdfm_cpu's one-pass Adam (csrc/cpu), or None when switched off with DFM_CPU_NATIVE=0."""
    global _cpu_lib
    import os
    if os.environ.get("DFM_CPU_NATIVE", "1") == "0":
        return None
    if _cpu_lib is None:
        import ctypes as C
        from .. import _native
        lib = _native.cpu()
        p, f = C.c_void_p, C.c_float
        lib.dfcpu_adam_dense.argtypes = [p, p, p, p, C.c_int64, f, f, f, f, C.c_int]
        lib.dfcpu_adam_dense.restype = None
        lib.dfcpu_adam_rows.argtypes = [p, p, p, p, p, C.c_int64, C.c_int64, f, f, f, f, C.c_int]
        lib.dfcpu_adam_rows.restype = None
        _cpu_lib = lib
    return _cpu_lib


def _host_f32(*ts) -> bool:
    return all(t is not None and not t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() for t in ts)


def canonical_optimizer(name: str) -> str:
    """Accepts the reference's spellings ('Adam', 'Adagrad', 'Momentum', 'ftrl', 'GD') in any case."""
    n = name.strip().lower()
    if n in ("sgd", "gradientdescent"):
        n = "gd"
    if n not in OPT_IDS:
        raise ValueError(f"unknown optimizer {name!r}; choose from Adam, Adagrad, Momentum, ftrl, GD")
    return n


class TFOptimizer:
    def __init__(self, kind: str, lr: float, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8,
                 momentum: float = 0.95, ftrl_lr_power: float = -0.5, ftrl_l1: float = 0.0, ftrl_l2: float = 0.0):
        self.kind = canonical_optimizer(kind)
        self.lr = float(lr)
        self.beta1, self.beta2, self.eps = beta1, beta2, eps
        self.momentum = momentum
        self.ftrl_lr_power, self.ftrl_l1, self.ftrl_l2 = ftrl_lr_power, ftrl_l1, ftrl_l2
        self.step = 0  # completed steps; the update in flight uses t = step + 1
        self.slots: Dict[str, torch.Tensor] = {}

    # ---------------------------------------------------------------- slots
    def slot_names(self, var: str) -> Tuple[Optional[str], Optional[str]]:
        a, b = _SLOT_NAMES[self.kind]
        return (f"{var}/{a}" if a else None, f"{var}/{b}" if b else None)

    def slot_init_values(self) -> Tuple[float, float]:
        return _SLOT_INIT[self.kind]

    def ensure_slots(self, var: str, p: torch.Tensor):
        n0, n1 = self.slot_names(var)
        i0, i1 = self.slot_init_values()
        if n0 and n0 not in self.slots:
            self.slots[n0] = torch.full_like(p, i0)
        if n1 and n1 not in self.slots:
            self.slots[n1] = torch.full_like(p, i1)
        return (self.slots.get(n0) if n0 else None, self.slots.get(n1) if n1 else None)

    # ---------------------------------------------------------------- update
    def _lr_t(self) -> float:
        if self.kind != "adam":
            return self.lr
        t = self.step + 1
        return self.lr * math.sqrt(1.0 - self.beta2 ** t) / (1.0 - self.beta1 ** t)

    def _apply(self, p: torch.Tensor, g: torch.Tensor, s0, s1) -> None:
        k = self.kind
        if k == "adam":
            s0.mul_(self.beta1).add_(g, alpha=1 - self.beta1)
            s1.mul_(self.beta2).addcmul_(g, g, value=1 - self.beta2)
            p.addcdiv_(s0, s1.sqrt().add_(self.eps), value=-self._lr_t())
        elif k == "adagrad":
            s0.addcmul_(g, g)
            p.addcdiv_(g, s0.sqrt(), value=-self.lr)
        elif k == "momentum":
            s0.mul_(self.momentum).add_(g)
            p.add_(s0, alpha=-self.lr)
        elif k == "ftrl":
            acc_new = s0 + g * g
            pw = -self.ftrl_lr_power
            sigma = (acc_new.pow(pw) - s0.pow(pw)) / self.lr
            s1.add_(g - sigma * p)
            quad = acc_new.pow(pw) / self.lr + 2 * self.ftrl_l2
            newp = torch.where(s1.abs() > self.ftrl_l1, (torch.sign(s1) * self.ftrl_l1 - s1) / quad,
                               torch.zeros_like(p))
            p.copy_(newp)
            s0.copy_(acc_new)
        else:
            p.add_(g, alpha=-self.lr)

    @torch.no_grad()
    def apply_dense(self, var: str, p: torch.Tensor, g: torch.Tensor) -> None:
        s0, s1 = self.ensure_slots(var, p)
        if self.kind == "adam" and _host_f32(p, s0, s1) and g.shape == p.shape and not g.is_cuda:
            lib = _native_adam()
            if lib is not None:
                g = g.detach().to(torch.float32).contiguous()
                lib.dfcpu_adam_dense(p.data_ptr(), s0.data_ptr(), s1.data_ptr(), g.data_ptr(), p.numel(),
                                     self._lr_t(), self.beta1, self.beta2, self.eps, torch.get_num_threads())
                return
        self._apply(p, g, s0, s1)

    @torch.no_grad()
    def apply_rows(self, var: str, p: torch.Tensor, rows: torch.Tensor, g_rows: torch.Tensor) -> None:
        """Lazy sparse update: ``rows`` are distinct indices into dim 0 of ``p``."""
        s0, s1 = self.ensure_slots(var, p)
        rows = rows.long()
        if self.kind == "adam" and _host_f32(p, s0, s1) and not g_rows.is_cuda and not rows.is_cuda \
                and g_rows.shape == (rows.numel(),) + tuple(p.shape[1:]):
            lib = _native_adam()
            if lib is not None:
                g = g_rows.detach().to(torch.float32).contiguous()
                rows = rows.contiguous()
                K = p[0].numel() if p.dim() > 1 else 1
                lib.dfcpu_adam_rows(p.data_ptr(), s0.data_ptr(), s1.data_ptr(), rows.data_ptr(), g.data_ptr(),
                                    rows.numel(), K, self._lr_t(), self.beta1, self.beta2, self.eps,
                                    torch.get_num_threads())
                return
        pr = p.index_select(0, rows)
        r0 = s0.index_select(0, rows) if s0 is not None else None
        r1 = s1.index_select(0, rows) if s1 is not None else None
        self._apply(pr, g_rows, r0, r1)
        p.index_copy_(0, rows, pr)
        if s0 is not None:
            s0.index_copy_(0, rows, r0)
        if s1 is not None:
            s1.index_copy_(0, rows, r1)

    def finish_step(self) -> None:
        self.step += 1

    # ------------------------------------------------------------ checkpoint
    def state_tensors(self) -> Dict[str, torch.Tensor]:
        """Slots (+ Adam's beta powers, as TF stores them) under TF checkpoint names."""
        d = dict(self.slots)
        if self.kind == "adam":
            # after `step` completed updates TF holds beta^(step+1)
            d["beta1_power"] = torch.tensor(self.beta1 ** (self.step + 1), dtype=torch.float32)
            d["beta2_power"] = torch.tensor(self.beta2 ** (self.step + 1), dtype=torch.float32)
        return d

    def load_state_tensors(self, d: Dict[str, torch.Tensor], step: int) -> None:
        self.step = int(step)
        for k in list(self.slots.keys()):
            if k in d:
                self.slots[k].copy_(torch.as_tensor(d[k]).to(self.slots[k].device))
