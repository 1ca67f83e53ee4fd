"""Python face of the native step engine (csrc/cuda/engine.cu).

``FusedEngine`` owns the parameters and optimizer slots as torch CUDA tensors, hands their
pointers to libdfm_cuda.so, and drives one fused training step (7 kernel launches: embed+FM
forward, tcgen05 MLP forward, tcgen05 backward chain, embedding-gradient GEMM+scatter,
split-K weight gradients, dense optimizer, sparse lazy optimizer) -- the work of one
``session.run(train_op)`` in the reference loop (HVD.py:430-435).

Dense parameters live in ONE flat fp32 vector ``[P_0 | ... | P_{L-1} | P_out | fm_bias]`` with
``P_l = [K_l + 1, N_l]`` (weights, then the bias as last row); TF-named views of it are
exposed by :meth:`variables` so checkpoints keep the reference's names and shapes.  Layer widths
that are not multiples of 16 are zero-padded internally (padding units have zero weights, zero
bias, zero gradient, so they stay exactly zero under every optimizer here).
"""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict
from typing import Dict, Optional

import torch

from .. import _native
from ..models.deepfm import DeepFM, DeepFMConfig
from ..utils.optimizers import OPT_IDS, _SLOT_INIT, _SLOT_NAMES, canonical_optimizer

MAX_LAYERS = 6


class _EngineConfig(C.Structure):
    _fields_ = [
        ("batch_max", C.c_int32), ("field_size", C.c_int32), ("embedding_size", C.c_int32), ("n_layers", C.c_int32),
        ("layers", C.c_int32 * MAX_LAYERS), ("keep_prob", C.c_float * MAX_LAYERS),
        ("feature_size", C.c_int64), ("optimizer", C.c_int32), ("table_update_mode", C.c_int32),
        ("lr", C.c_float), ("l2_reg", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
        ("momentum", C.c_float), ("ftrl_lr_power", C.c_float), ("ftrl_l1", C.c_float), ("ftrl_l2", C.c_float),
        ("seed", C.c_uint64),
        ("batch_norm", C.c_int32), ("loss_type", C.c_int32), ("bn_decay", C.c_float), ("bn_eps", C.c_float),
    ]


class _EngineTensors(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "fm_w", "fm_w_s0", "fm_w_s1", "fm_v", "fm_v_s0", "fm_v_s1", "dense", "dense_s0", "dense_s1", "bn_moving")]


class _DevMem:
    """Zero-copy window onto engine-owned device memory (via __cuda_array_interface__)."""

    def __init__(self, ptr: int, n: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def _round_up(a: int, b: int) -> int:
    return (a + b - 1) // b * b


def fused_engine_supported(cfg: DeepFMConfig) -> Optional[str]:
    """None if the fused kernels cover this configuration, else the reason they do not."""
    if cfg.embedding_size % 16 or cfg.embedding_size > 64:
        return "embedding_size must be a multiple of 16 and <= 64"
    if not 1 <= len(cfg.deep_layers) <= MAX_LAYERS:
        return f"1..{MAX_LAYERS} deep layers"
    if max(cfg.deep_layers) > 512:
        return "deep layer widths must be <= 512"
    return None


class FusedEngine:
    def __init__(self, cfg: DeepFMConfig, batch_max: int, optimizer: str = "Adam", lr: float = 5e-4,
                 table_update_mode: str = "lazy", seed: int = 0, device: Optional[torch.device] = None,
                 beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, momentum: float = 0.95,
                 ftrl_lr_power: float = -0.5, ftrl_l1: float = 0.0, ftrl_l2: float = 0.0, provider=None):
        """``provider`` (parallel.symmetric): allocates the embedding tables where every rank can address them,
        which the multi-GPU path needs (the owner of a row stores its new value into every replica)."""
        why = fused_engine_supported(cfg)
        if why:
            raise ValueError(f"fused engine does not cover this configuration: {why}")
        if not torch.cuda.is_available():
            raise RuntimeError("FusedEngine needs a CUDA device (sm_100a)")
        self.cfg = cfg
        self.device = torch.device(device or f"cuda:{torch.cuda.current_device()}")
        self._dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.batch_max = int(batch_max)
        self.opt_kind = canonical_optimizer(optimizer)
        self.lr = float(lr)
        self.table_update_mode = table_update_mode
        self.beta1, self.beta2 = float(beta1), float(beta2)
        self._lib = _native.cuda()
        self._type_lib()

        L = len(cfg.deep_layers)
        self.L = L
        self.Np = [_round_up(n, 16) for n in cfg.deep_layers]           # padded widths
        self.K = [cfg.field_size * cfg.embedding_size] + self.Np        # K[l] input width of layer l
        self.N = self.Np + [1]
        # flat dense vector: [P_0 | .. | P_{L-1} | (batch norm: [gamma_l | beta_l] per layer) | P_out | fm_bias]
        self.off, self.off_bn, self.off_mov = [], [], []
        o = om = 0
        for l in range(L):
            self.off.append(o)
            o += (self.K[l] + 1) * self.N[l]
        if cfg.batch_norm:
            for l in range(L):
                self.off_bn.append(o)
                o += 2 * self.N[l]
                self.off_mov.append(om)
                om += 2 * self.N[l]
        self.off.append(o)            # P_out
        o += (self.K[L] + 1) * self.N[L]
        self.off.append(o)            # fm_bias
        self.n_dense = o + 1

        dev, f32 = self.device, torch.float32
        V, Ke = cfg.feature_size, cfg.embedding_size
        i0, i1 = _SLOT_INIT[self.opt_kind]
        self.world, self.rank = 1, 0
        self.grad_table = None
        if provider is not None:
            self.fm_w = provider.alloc((V,), f32, dev)
            self.fm_v = provider.alloc((V, Ke), f32, dev)
        else:
            self.fm_w = torch.zeros(V, device=dev, dtype=f32)
            self.fm_v = torch.zeros(V, Ke, device=dev, dtype=f32)
        self.dense = torch.zeros(self.n_dense, device=dev, dtype=f32)
        self.fm_w_s0, self.fm_w_s1 = torch.full_like(self.fm_w, i0), torch.full_like(self.fm_w, i1)
        self.fm_v_s0, self.fm_v_s1 = torch.full_like(self.fm_v, i0), torch.full_like(self.fm_v, i1)
        self.dense_s0, self.dense_s1 = torch.full_like(self.dense, i0), torch.full_like(self.dense, i1)
        self.bn_moving = None
        if cfg.batch_norm:
            self.bn_moving = torch.zeros(max(om, 1), device=dev, dtype=f32)
            for l in range(L):        # gamma = 1, moving variance = 1 (padding units included: they only ever see zeros)
                self.dense[self.off_bn[l]:self.off_bn[l] + self.N[l]] = 1.0
                self.bn_moving[self.off_mov[l] + self.N[l]:self.off_mov[l] + 2 * self.N[l]] = 1.0

        ec = _EngineConfig()
        ec.batch_max, ec.field_size, ec.embedding_size, ec.n_layers = self.batch_max, cfg.field_size, Ke, L
        for l in range(L):
            ec.layers[l] = self.Np[l]
            ec.keep_prob[l] = float(cfg.dropout_keep[l])
        ec.feature_size = V
        ec.optimizer = OPT_IDS[self.opt_kind]
        ec.table_update_mode = {"lazy": 0, "dense_exact": 1}[table_update_mode]
        ec.lr, ec.l2_reg = self.lr, cfg.l2_reg
        ec.beta1, ec.beta2, ec.eps, ec.momentum = beta1, beta2, eps, momentum
        ec.ftrl_lr_power, ec.ftrl_l1, ec.ftrl_l2 = ftrl_lr_power, ftrl_l1, ftrl_l2
        ec.seed = seed & 0xFFFFFFFFFFFFFFFF
        ec.batch_norm = int(bool(cfg.batch_norm))
        ec.loss_type = {"log_loss": 0, "square_loss": 1}[cfg.loss_type]
        ec.bn_decay = float(cfg.batch_norm_decay)
        from ..models.deepfm import BN_EPS
        ec.bn_eps = BN_EPS
        et = _EngineTensors(*([t.data_ptr() for t in (
            self.fm_w, self.fm_w_s0, self.fm_w_s1, self.fm_v, self.fm_v_s0, self.fm_v_s1,
            self.dense, self.dense_s0, self.dense_s1)] + [self.bn_moving.data_ptr() if cfg.batch_norm else None]))
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            _native.check(self._lib.dfm_engine_create(C.byref(ec), C.byref(et), C.byref(h)), "dfm_engine_create")
        self._h = h
        self._bufs: Dict[str, torch.Tensor] = {}
        self.state_t = self.buffer("state", dtype="i4")   # raw int32 view of StepState

    # ------------------------------------------------------------------ plumbing
    def _type_lib(self):
        lib = self._lib
        if getattr(lib, "_engine_typed", False):
            return
        lib.dfm_engine_create.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
        lib.dfm_engine_destroy.argtypes = [C.c_void_p]
        lib.dfm_engine_set_state.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_int, C.c_void_p]
        lib.dfm_engine_sync_weights.argtypes = [C.c_void_p, C.c_void_p]
        lib.dfm_engine_forward.argtypes = [C.c_void_p] * 5
        lib.dfm_engine_train_step.argtypes = [C.c_void_p] * 5
        lib.dfm_engine_buffer.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_longlong)]
        lib.dfm_engine_launches_per_step.argtypes = [C.c_void_p]
        lib.dfm_auc_hist.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        lib.dfm_engine_slab_layout.argtypes = [C.c_void_p] + [C.POINTER(C.c_longlong)] * 3
        lib.dfm_engine_capture.argtypes = [C.c_void_p, C.c_void_p]
        lib.dfm_engine_slab_ptr.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        lib.dfm_engine_replay.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        lib.dfm_engine_step_pipelined.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.c_void_p]
        lib.dfm_engine_reset_pipeline.argtypes = [C.c_void_p]
        lib.dfm_engine_loss_ring.argtypes = [C.c_void_p, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_int)]
        lib._engine_typed = True

    def current_stream(self) -> "torch.cuda.Stream":
        """``torch.cuda.current_stream(self.device)`` without its per-call Python cost (index normalisation, a new
        Stream object: ~6 us, three times per step -- a third of the main thread in the file-to-loss pipeline).  The
        builtin underneath answers in 0.2 us; the Stream object and its raw handle are rebuilt only when that answer
        changes."""
        get = getattr(torch._C, "_cuda_getCurrentStream", None)
        if get is None:                                   # private API gone: the slow, public way
            self._cur_stream = torch.cuda.current_stream(self.device)
            self._cur_stream_ptr = C.c_void_p(self._cur_stream.cuda_stream)
            return self._cur_stream
        key = get(self._dev_index)
        if key != getattr(self, "_cur_stream_key", None):
            self._cur_stream = torch.cuda.current_stream(self.device)
            self._cur_stream_ptr = C.c_void_p(self._cur_stream.cuda_stream)
            self._cur_stream_key = key
        return self._cur_stream

    def _stream(self) -> C.c_void_p:
        self.current_stream()
        return self._cur_stream_ptr

    def buffer(self, name: str, index: int = 0, dtype: str = "f4") -> torch.Tensor:
        """Zero-copy torch view of an engine buffer (dtype: 'f4', 'i4', or 'bf16')."""
        key = f"{name}:{index}:{dtype}"
        if key not in self._bufs:
            p, n = C.c_void_p(), C.c_longlong()
            _native.check(self._lib.dfm_engine_buffer(self._h, name.encode(), index, C.byref(p), C.byref(n)),
                          "dfm_engine_buffer")
            typestr = {"f4": "<f4", "i4": "<i4", "bf16": "<i2"}[dtype]
            t = torch.as_tensor(_DevMem(p.value, n.value, typestr), device=self.device)
            if dtype == "bf16":
                t = t.view(torch.bfloat16)
            self._bufs[key] = t
        return self._bufs[key]

    def close(self):
        if getattr(self, "_h", None):
            self._lib.dfm_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ parameters
    def _layer_views(self, flat: torch.Tensor, l: int):
        """(weights [K_l, N_l], bias [N_l]) views of layer l inside a flat dense vector."""
        K, N, o = self.K[l], self.N[l], self.off[l]
        return flat[o:o + K * N].view(K, N), flat[o + K * N:o + (K + 1) * N]

    @torch.no_grad()
    def load_from_module(self, m: DeepFM) -> None:
        cfg = self.cfg
        self.fm_w.copy_(m.fm_w.detach())
        self.fm_v.copy_(m.fm_v.detach())
        self.dense.zero_()
        k_prev = cfg.field_size * cfg.embedding_size
        for l in range(self.L):
            w, b = self._layer_views(self.dense, l)
            n = cfg.deep_layers[l]
            w[:k_prev, :n] = m.mlp_w[l].detach()
            b[:n] = m.mlp_b[l].detach()
            if cfg.batch_norm:
                ob, om, N = self.off_bn[l], self.off_mov[l], self.N[l]
                self.dense[ob:ob + n] = m.bn_gamma[l].detach()
                self.dense[ob + N:ob + N + n] = m.bn_beta[l].detach()
                self.bn_moving[om:om + n] = getattr(m, f"bn_mean_{l}")
                self.bn_moving[om + N:om + N + n] = getattr(m, f"bn_var_{l}")
            k_prev = n
        w, b = self._layer_views(self.dense, self.L)
        w[:k_prev, :] = m.out_w.detach()
        b[:] = m.out_b.detach()
        self.dense[self.off[self.L + 1]] = m.fm_bias.detach()[0]
        self.sync_weights()

    @torch.no_grad()
    def store_to_module(self, m: DeepFM) -> None:
        m.load_tf_variables({k: v for k, v in self.variables().items()}, strict=True)

    def variables(self) -> "OrderedDict[str, torch.Tensor]":
        """TF-named parameter tensors (unpadded views / slices of the engine's storage)."""
        return self._named(self.fm_w, self.fm_v, self.dense, self.bn_moving)

    def _named(self, fm_w, fm_v, dense, moving=None) -> "OrderedDict[str, torch.Tensor]":
        cfg = self.cfg
        d: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        d["fm_bias"] = dense[self.off[self.L + 1]:self.off[self.L + 1] + 1]
        d["fm_w"] = fm_w
        d["fm_v"] = fm_v
        k_prev = cfg.field_size * cfg.embedding_size
        for l in range(self.L):
            w, b = self._layer_views(dense, l)
            n = cfg.deep_layers[l]
            d[f"Deep-part/mlp{l}/weights"] = w[:k_prev, :n]
            d[f"Deep-part/mlp{l}/biases"] = b[:n]
            if cfg.batch_norm:
                ob, N = self.off_bn[l], self.N[l]
                d[f"Deep-part/bn_{l}/beta"] = dense[ob + N:ob + N + n]
                d[f"Deep-part/bn_{l}/gamma"] = dense[ob:ob + n]
                if moving is not None:
                    om = self.off_mov[l]
                    d[f"Deep-part/bn_{l}/moving_mean"] = moving[om:om + n]
                    d[f"Deep-part/bn_{l}/moving_variance"] = moving[om + N:om + N + n]
            k_prev = n
        w, b = self._layer_views(dense, self.L)
        d["Deep-part/deep_out/weights"] = w[:k_prev, :]
        d["Deep-part/deep_out/biases"] = b
        return d

    def slot_variables(self) -> "OrderedDict[str, torch.Tensor]":
        """Optimizer slots under TF checkpoint names (``<var>/Adam`` ...)."""
        n0, n1 = _SLOT_NAMES[self.opt_kind]
        out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        for suffix, (w, v, dn) in ((n0, (self.fm_w_s0, self.fm_v_s0, self.dense_s0)),
                                   (n1, (self.fm_w_s1, self.fm_v_s1, self.dense_s1))):
            if suffix is None:
                continue
            for k, t in self._named(w, v, dn).items():
                out[f"{k}/{suffix}"] = t
        return out

    def owned_row_mask(self) -> Optional[torch.Tensor]:
        """Multi-GPU lazy update: bool [V], True where this rank holds the current optimizer slots of the table row
        (row id belongs to rank id % world; csrc/cuda/comm.cu).  None when every row's slots are local."""
        if self.world <= 1 or self.table_update_mode != "lazy":
            return None
        return (torch.arange(self.cfg.feature_size, device=self.device) % self.world) == self.rank

    def sync_weights(self) -> None:
        """Call after writing to ``dense`` from Python (init, restore, broadcast)."""
        _native.check(self._lib.dfm_engine_sync_weights(self._h, self._stream()), "dfm_engine_sync_weights")

    # ------------------------------------------------------------------ stepping
    def set_state(self, count: int, train: bool, lr: Optional[float] = None, step: int = -1) -> None:
        _native.check(self._lib.dfm_engine_set_state(self._h, int(count), float(self.lr if lr is None else lr),
                                                     int(train), int(step), self._stream()), "dfm_engine_set_state")

    def _check_inputs(self, ids, vals, labels):
        assert ids.is_cuda and ids.dtype == torch.int32 and ids.is_contiguous(), "feat_ids must be contiguous int32 CUDA"
        assert vals.is_cuda and vals.dtype == torch.float32 and vals.is_contiguous()
        assert ids.shape == vals.shape and ids.shape[1] == self.cfg.field_size and ids.shape[0] <= self.batch_max
        if labels is not None:
            assert labels.is_cuda and labels.dtype == torch.float32 and labels.numel() == ids.shape[0]

    def forward(self, ids: torch.Tensor, vals: torch.Tensor, labels: Optional[torch.Tensor] = None,
                set_state: bool = True) -> torch.Tensor:
        """Inference forward; returns the probability buffer view ``[count]`` (valid until the next call)."""
        self._check_inputs(ids, vals, labels)
        n = ids.shape[0]
        if set_state:
            self.set_state(n, train=False)
        _native.check(self._lib.dfm_engine_forward(self._h, ids.data_ptr(), vals.data_ptr(),
                                                   labels.data_ptr() if labels is not None else None,
                                                   self._stream()), "dfm_engine_forward")
        return self.buffer("prob")[:n]

    def train_step(self, ids: torch.Tensor, vals: torch.Tensor, labels: torch.Tensor, set_state: bool = True) -> None:
        """One optimizer step on the batch.  Asynchronous; read :meth:`loss` afterwards if needed."""
        self._check_inputs(ids, vals, labels)
        if set_state:
            self.set_state(ids.shape[0], train=True)
        _native.check(self._lib.dfm_engine_train_step(self._h, ids.data_ptr(), vals.data_ptr(), labels.data_ptr(),
                                                      self._stream()), "dfm_engine_train_step")

    # StepState field order: count, step, lr, loss_sum, n_unique, bad_ids, train, pad
    def loss_sum_tensor(self) -> torch.Tensor:
        return self.state_t.view(torch.float32)[3:4]

    def step_tensor(self) -> torch.Tensor:
        return self.state_t[1:2]

    def read_state(self) -> Dict[str, float]:
        s = self.state_t.cpu()
        f = s.view(torch.float32)
        return {"count": int(s[0]), "step": int(s[1]), "lr": float(f[2]), "loss_sum": float(f[3]),
                "n_unique": int(s[4]), "bad_ids": int(s[5]), "train": int(s[6])}

    # ------------------------------------------------------------------ graph replay / host pipeline
    def capture(self, count: Optional[int] = None) -> None:
        """Records the train step as a CUDA graph (one per device input slab).  Fixes ``count`` and the
        train flag; the optimizer step counter keeps advancing on the device."""
        self.set_state(self.batch_max if count is None else count, train=True)
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            _native.check(self._lib.dfm_engine_capture(self._h, C.c_void_p(side.cuda_stream)), "dfm_engine_capture")
        torch.cuda.current_stream(self.device).wait_stream(side)
        self._captured = True

    def device_slab(self, which: int):
        """(slab_bytes_tensor, ids, vals, labels) views of device input slab ``which`` (after capture)."""
        p = C.c_void_p()
        _native.check(self._lib.dfm_engine_slab_ptr(self._h, which, C.byref(p)), "dfm_engine_slab_ptr")
        tot, ov, ol = C.c_longlong(), C.c_longlong(), C.c_longlong()
        self._lib.dfm_engine_slab_layout(self._h, C.byref(tot), C.byref(ov), C.byref(ol))
        slab = torch.as_tensor(_DevMem(p.value, tot.value, "|u1"), device=self.device)
        B, Fd = self.batch_max, self.cfg.field_size
        return (slab, slab[:ov.value].view(torch.int32).view(B, Fd), slab[ov.value:ol.value].view(torch.float32).view(B, Fd),
                slab[ol.value:].view(torch.float32))

    def replay(self, which: int = 0) -> None:
        _native.check(self._lib.dfm_engine_replay(self._h, which, self._stream()), "dfm_engine_replay")

    def step_pipelined(self, cur: torch.Tensor, nxt: Optional[torch.Tensor] = None) -> int:
        """Trains on slab ``cur`` while the graph's prefetch node copies ``nxt`` (pinned host or device slab)
        into the other input slab; pass that tensor as ``cur`` next time.  Returns the loss-ring slot."""
        slot = C.c_int(0)
        _native.check(self._lib.dfm_engine_step_pipelined(
            self._h, C.c_void_p(cur.data_ptr()), C.c_void_p(nxt.data_ptr()) if nxt is not None else None,
            C.byref(slot), self._stream()), "dfm_engine_step_pipelined")
        return slot.value

    def reset_pipeline(self) -> None:
        """Forget what the prefetch node staged (call when slab addresses may be reused with new contents)."""
        _native.check(self._lib.dfm_engine_reset_pipeline(self._h), "dfm_engine_reset_pipeline")

    def loss_ring(self):
        """ctypes float array over the engine's pinned loss ring (one entry per input slab)."""
        ptr, n = C.POINTER(C.c_float)(), C.c_int(0)
        _native.check(self._lib.dfm_engine_loss_ring(self._h, C.byref(ptr), C.byref(n)), "dfm_engine_loss_ring")
        return (C.c_float * n.value).from_address(C.addressof(ptr.contents))   # a view, not a copy

    def launches_per_step(self) -> int:
        return int(self._lib.dfm_engine_launches_per_step(self._h))


class EmulatedRanks:
    """N fused engines on ONE GPU acting as the N ranks of a synchronous data-parallel job.

    Each engine runs its own forward / backward; the two collective kernels (dense all-reduce + optimizer, sparse
    owner-sharded exchange) are launched once, cooperatively, over all virtual ranks' data
    (``dfm_engine_emu_collectives``).  This is how the multi-rank protocol of csrc/cuda/comm.cu is tested and
    profiled on a box with a single GPU; real jobs use one process per GPU (parallel/symmetric.py)."""

    def __init__(self, cfg: DeepFMConfig, n: int, batch_max: int, **engine_kw):
        from ..parallel.symmetric import LocalProvider
        self.n = n
        self.engines = [FusedEngine(cfg, batch_max, provider=LocalProvider(), **engine_kw) for _ in range(n)]
        lib = self.engines[0]._lib
        lib.dfm_engine_emu_attach.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        lib.dfm_engine_emu_collectives.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        self._lib = lib
        self._handles = (C.c_void_p * n)(*[e._h for e in self.engines])
        grads = None
        if self.engines[0].table_update_mode == "dense_exact":
            for e in self.engines:
                e.grad_table = torch.zeros(cfg.feature_size, cfg.embedding_size + 8, device=e.device)
            grads = (C.c_void_p * n)(*[C.c_void_p(e.grad_table.data_ptr()) for e in self.engines])
        _native.check(lib.dfm_engine_emu_attach(self._handles, n, grads), "dfm_engine_emu_attach")
        for r, e in enumerate(self.engines):
            e.world, e.rank = n, r
            e._bufs.clear()
            e.state_t = e.buffer("state", dtype="i4")

    def load_from_module(self, m: DeepFM) -> None:
        for e in self.engines:
            e.load_from_module(m)

    def train_step(self, batches) -> None:
        """``batches[r] = (ids int32 [B,F], vals, labels)`` of virtual rank r; one synchronous step of all ranks."""
        for e, (ids, vals, labels) in zip(self.engines, batches):
            e.train_step(ids, vals, labels)
        _native.check(self._lib.dfm_engine_emu_collectives(self._handles, self.n, self.engines[0]._stream()),
                      "dfm_engine_emu_collectives")

    def close(self) -> None:
        torch.cuda.synchronize()
        for e in self.engines:
            e.close()


def auc_histogram(prob: torch.Tensor, label: torch.Tensor, hist: torch.Tensor, num_thresholds: int = 200) -> None:
    """Accumulates tf.metrics.auc-style confusion buckets on the GPU (csrc/cuda/optim.cu)."""
    assert prob.is_cuda and hist.dtype == torch.int64 and hist.numel() == 2 * (num_thresholds + 1)
    lib = _native.cuda()
    lib.dfm_auc_hist.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
    _native.check(lib.dfm_auc_hist(prob.data_ptr(), label.data_ptr(), prob.numel(), hist.data_ptr(), num_thresholds,
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream)), "dfm_auc_hist")
