"""Oracle model math, TF-formula optimizers, AUC, export -- all on CPU."""
import numpy as np
import pytest
import torch

from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.utils.metrics import StreamingAUC, exact_auc, tf_auc_numpy
from deepfm_b200.utils.optimizers import TFOptimizer, canonical_optimizer


def test_forward_matches_hand_written_numpy():
    cfg = DeepFMConfig(field_size=5, feature_size=50, embedding_size=4, deep_layers=[8, 4], dropout_keep=[1, 1])
    m = DeepFM(cfg, seed=3)
    with torch.no_grad():
        for p in m.parameters():
            p.copy_(torch.randn_like(p) * 0.3)
    ids = torch.tensor([[1, 2, 3, 4, 4], [9, 8, 7, 6, 5]])
    vals = torch.rand(2, 5)
    got = m(ids, vals).detach().numpy()
    W = {k: v.detach().numpy() for k, v in m.tf_variables().items()}
    x, i = vals.numpy(), ids.numpy()
    y_w = (W["fm_w"][i] * x).sum(1)
    e = W["fm_v"][i] * x[..., None]
    y_v = 0.5 * ((e.sum(1) ** 2) - (e ** 2).sum(1)).sum(1)
    h = e.reshape(2, -1)
    for l in range(2):
        h = np.maximum(h @ W[f"Deep-part/mlp{l}/weights"] + W[f"Deep-part/mlp{l}/biases"], 0)
    y = W["fm_bias"] + y_w + y_v + (h @ W["Deep-part/deep_out/weights"]).reshape(-1) + W["Deep-part/deep_out/biases"]
    np.testing.assert_allclose(got, y, rtol=1e-5, atol=1e-5)
    labels = torch.tensor([1.0, 0.0])
    loss = m.loss(m(ids, vals), labels).item()
    bce = np.mean(np.maximum(y, 0) - y * labels.numpy() + np.log1p(np.exp(-np.abs(y))))
    reg = cfg.l2_reg * 0.5 * ((W["fm_w"] ** 2).sum() + (W["fm_v"] ** 2).sum())
    assert abs(loss - (bce + reg)) < 1e-5


def test_variable_names_and_shapes_follow_the_reference_graph():
    cfg = DeepFMConfig(39, 117581, 32, [256, 128, 64], [0.5] * 3, batch_norm=True)
    m = DeepFM(cfg, seed=0)
    v = m.tf_variables()
    assert tuple(v["fm_bias"].shape) == (1,) and tuple(v["fm_w"].shape) == (117581,)
    assert tuple(v["fm_v"].shape) == (117581, 32)
    assert tuple(v["Deep-part/mlp0/weights"].shape) == (1248, 256)
    assert tuple(v["Deep-part/mlp2/biases"].shape) == (64,)
    assert tuple(v["Deep-part/deep_out/weights"].shape) == (64, 1)
    for leaf in ("beta", "gamma", "moving_mean", "moving_variance"):
        assert tuple(v[f"Deep-part/bn_1/{leaf}"].shape) == (128,)
    assert DeepFM(DeepFMConfig(39, 117581), seed=0).num_parameters() == 4241135
    # initialisers: glorot normal on the FM tables, zeros on biases
    assert abs(m.fm_v.std().item() - (2 / (117581 + 32)) ** 0.5) < 3e-4 and m.fm_bias.item() == 0


def test_dropout_is_keep_probability_and_bn_follows_relu():
    cfg = DeepFMConfig(4, 20, 4, [64], [0.25], batch_norm=True, batch_norm_decay=0.9)
    m = DeepFM(cfg, seed=0)
    ids, vals = torch.randint(0, 20, (512, 4)), torch.rand(512, 4)
    before = m.bn_mean_0.clone()
    m(ids, vals, training=False)
    assert torch.equal(m.bn_mean_0, before)                       # eval never touches the moving stats (Q5)
    m(ids, vals, training=True)
    assert not torch.equal(m.bn_mean_0, before)
    y_w, y_v, e = m.fm_terms(ids, vals)
    h = torch.relu(e.reshape(512, -1) @ m.mlp_w[0] + m.mlp_b[0])
    assert torch.allclose(m.bn_mean_0, 0.1 * h.mean(0), atol=1e-6)    # decay 0.9 from zero
    torch.manual_seed(0)
    cfg2 = DeepFMConfig(4, 20, 4, [4096], [0.25])
    m2 = DeepFM(cfg2, seed=0)
    with torch.no_grad():
        m2.mlp_b[0].fill_(1.0)
    e = m2.fm_terms(ids[:8], vals[:8])[2].reshape(8, -1)
    h = torch.relu(e @ m2.mlp_w[0] + m2.mlp_b[0])
    hd = torch.nn.functional.dropout(h, p=0.75, training=True)
    assert 0.22 < (hd != 0).float().mean().item() / (h != 0).float().mean().item() < 0.28


def _np_step(kind, p, g, s0, s1, lr, t):
    if kind == "adam":
        b1, b2, eps = 0.9, 0.999, 1e-8
        s0 = b1 * s0 + (1 - b1) * g
        s1 = b2 * s1 + (1 - b2) * g * g
        lr_t = lr * np.sqrt(1 - b2 ** t) / (1 - b1 ** t)
        p = p - lr_t * s0 / (np.sqrt(s1) + eps)
    elif kind == "adagrad":
        s0 = s0 + g * g
        p = p - lr * g / np.sqrt(s0)
    elif kind == "momentum":
        s0 = 0.95 * s0 + g
        p = p - lr * s0
    elif kind == "ftrl":
        new = s0 + g * g
        sigma = (np.sqrt(new) - np.sqrt(s0)) / lr
        s1 = s1 + g - sigma * p
        p = -s1 / (np.sqrt(new) / lr)
        s0 = new
    else:
        p = p - lr * g
    return p, s0, s1


@pytest.mark.parametrize("kind", ["adam", "adagrad", "momentum", "ftrl", "gd"])
def test_optimizers_match_tf_update_formulas(kind):
    rng = np.random.default_rng(1)
    p0 = rng.standard_normal(64).astype(np.float64)
    opt = TFOptimizer(kind, lr=0.05)
    p = torch.tensor(p0.copy())
    i0, i1 = opt.slot_init_values()
    s0, s1 = np.full(64, i0), np.full(64, i1)
    pn = p0.copy()
    for t in range(1, 6):
        g = rng.standard_normal(64)
        opt.apply_dense("w", p, torch.tensor(g))
        opt.finish_step()
        pn, s0, s1 = _np_step(kind, pn, g, s0, s1, 0.05, t)
    np.testing.assert_allclose(p.numpy(), pn, rtol=1e-9, atol=1e-12)
    names = [n for n in opt.slot_names("w") if n]
    assert names == {"adam": ["w/Adam", "w/Adam_1"], "adagrad": ["w/Adagrad"], "momentum": ["w/Momentum"],
                     "ftrl": ["w/Ftrl", "w/Ftrl_1"], "gd": []}[kind]


def test_lazy_rows_only_touch_their_rows():
    opt = TFOptimizer("Adam", 0.1)
    p = torch.ones(10, 4)
    rows = torch.tensor([2, 7])
    opt.apply_rows("fm_v", p, rows, torch.ones(2, 4))
    assert (p[[0, 1, 3, 4, 5, 6, 8, 9]] == 1).all() and (p[rows] < 1).all()
    assert (opt.slots["fm_v/Adam"][[0, 1]] == 0).all() and (opt.slots["fm_v/Adam"][rows] != 0).all()
    assert canonical_optimizer("GD") == "gd" and canonical_optimizer("ftrl") == "ftrl"


def test_auc_modes():
    from sklearn.metrics import roc_auc_score
    rng = np.random.default_rng(0)
    p = rng.random(20000).astype(np.float32)
    y = (rng.random(20000) < p).astype(np.float32)
    assert abs(exact_auc(y, p) - roc_auc_score(y, p)) < 1e-12
    assert abs(tf_auc_numpy(y, p) - roc_auc_score(y, p)) < 2e-3       # 200-threshold trapezoid approximation
    s = StreamingAUC()
    for i in range(0, 20000, 1024):
        s.update(y[i:i + 1024], p[i:i + 1024])
    assert abs(s.result() - tf_auc_numpy(y, p)) < 1e-12 and abs(s.result_exact() - exact_auc(y, p)) < 1e-12
    # callers hand in views of reused ring buffers: what was seen must not change when the buffer does
    s2, buf_y, buf_p = StreamingAUC(), np.empty(1024, np.float32), np.empty(1024, np.float32)
    for i in range(0, 20000, 1024):
        n = min(1024, 20000 - i)
        buf_y[:n], buf_p[:n] = y[i:i + n], p[i:i + n]
        s2.update(buf_y[:n], buf_p[:n])
    assert abs(s2.result_exact() - exact_auc(y, p)) < 1e-12
    # ties and degenerate inputs
    assert exact_auc(np.array([0, 1, 0, 1.0]), np.array([0.5, 0.5, 0.5, 0.5])) == 0.5
    assert np.isnan(exact_auc(np.ones(4), np.random.rand(4)))


def test_export_and_serving_signature(tmp_path):
    from deepfm_b200.utils.export import export_servable, load_servable
    cfg = DeepFMConfig(39, 500, 8, [16, 8], [0.5, 0.5])
    m = DeepFM(cfg, seed=5)
    out = export_servable(m, str(tmp_path / "servable"), timestamp=1600000000)
    assert out.endswith("1600000000")
    assert export_servable(m, str(tmp_path / "servable"), timestamp=1600000000).endswith("1600000001")
    serve = load_servable(str(tmp_path / "servable"))
    ids = np.random.randint(0, 500, (7, 39)).astype(np.int64)
    vals = np.random.rand(7, 39).astype(np.float32)
    got = serve(ids, vals)
    assert set(got) == {"prob"} and got["prob"].shape == (7,) and got["prob"].dtype == np.float32
    ref = torch.sigmoid(m(torch.from_numpy(ids), torch.from_numpy(vals))).detach().numpy()
    np.testing.assert_allclose(got["prob"], ref, rtol=1e-6)


def test_exact_auc_torch_matches_numpy_with_ties():
    import numpy as np
    import torch
    from deepfm_b200.utils.metrics import exact_auc, exact_auc_torch
    rng = np.random.default_rng(0)
    preds = np.round(rng.random(5000), 2).astype(np.float32)        # many ties
    labels = (rng.random(5000) < 0.3).astype(np.float32)
    assert abs(exact_auc_torch(torch.from_numpy(labels), torch.from_numpy(preds)) - exact_auc(labels, preds)) < 1e-12


def test_autograd_gradients_match_finite_differences_in_fp64():
    """SURVEY 4 (unit - model math): the eager restatement of HVD.py:190-268 is the oracle of every kernel test, so its
    own gradients are checked once against central differences in double precision (log loss and square loss, with
    and without batch norm, L2 term included)."""
    import torch
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(0, 30, (6, 4), generator=g)
    ids[0, 1] = ids[0, 2]                                   # the same row twice in one example
    vals = torch.rand(6, 4, generator=g, dtype=torch.float64)
    labels = torch.tensor([1.0, 0.0, 0.0, 1.0, 0.0, 1.0], dtype=torch.float64)
    for loss_type, bn in (("log_loss", False), ("square_loss", False), ("log_loss", True)):
        cfg = DeepFMConfig(4, 30, 3, [5, 4], [1.0, 1.0], batch_norm=bn, l2_reg=1e-2, loss_type=loss_type)
        m = DeepFM(cfg, seed=2).double()
        with torch.no_grad():
            m.fm_v.mul_(3); m.fm_w.mul_(3)                  # larger than the init, but no saturated sigmoid
            for b in m.mlp_b:
                b.uniform_(0.1, 0.3)                       # keep the ReLUs away from their kink
        assert m(ids, vals).abs().max() < 6
        keys = [k for k, _ in m.named_parameters()]

        def f(*ps):
            state = dict(zip(keys, ps))
            for k, b in m.named_buffers():                  # training-mode BN moves its averages in place: give it copies
                state[k] = b.clone()
            logits = torch.func.functional_call(m, state, (ids, vals), {"training": True})
            reg = cfg.l2_reg * 0.5 * (state["fm_w"].pow(2).sum() + state["fm_v"].pow(2).sum())
            return m.data_loss(logits, labels) + reg

        inputs = tuple(p.detach().clone().requires_grad_(True) for _, p in m.named_parameters())
        assert torch.autograd.gradcheck(f, inputs, eps=1e-6, atol=1e-7, rtol=1e-4), (loss_type, bn)


def test_forward_rows_is_forward_with_deduplicated_gradients():
    """parallel/ps.py trains on the batch's distinct rows: same logits, and the gradient w.r.t. the distinct rows is the
    scatter-add of the per-position gradients (what IndexedSlices de-duplication gives)."""
    import torch
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    cfg = DeepFMConfig(6, 50, 4, [8, 4], [1.0, 1.0])
    m = DeepFM(cfg, seed=1)
    g = torch.Generator().manual_seed(3)
    ids = torch.randint(0, 12, (16, 6), generator=g)         # few distinct ids: many repeats
    vals = torch.rand(16, 6, generator=g)
    labels = (torch.rand(16, generator=g) < 0.5).float()
    m.data_loss(m(ids, vals), labels).backward()
    uniq, inverse = ids.unique(return_inverse=True)
    w_rows = m.fm_w.detach()[uniq].clone().requires_grad_(True)
    v_rows = m.fm_v.detach()[uniq].clone().requires_grad_(True)
    logits = m.forward_rows(w_rows, v_rows, inverse, vals)
    torch.testing.assert_close(logits, m(ids, vals), atol=1e-6, rtol=1e-6)
    m.data_loss(logits, labels).backward()
    torch.testing.assert_close(w_rows.grad, m.fm_w.grad[uniq], atol=1e-6, rtol=1e-5)
    torch.testing.assert_close(v_rows.grad, m.fm_v.grad[uniq], atol=1e-6, rtol=1e-5)
    untouched = torch.ones(50, dtype=torch.bool); untouched[uniq] = False
    assert m.fm_v.grad[untouched].abs().max() == 0


def test_fm_second_order_identity_equals_the_pairwise_sum():
    """0.5 * ((sum_f v_f x_f)^2 - sum_f (v_f x_f)^2) summed over K == sum_{f<g} <v_f, v_g> x_f x_g (HVD.py:200-206)."""
    import torch
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    m = DeepFM(DeepFMConfig(5, 40, 6, [4], [1.0]), seed=4).double()
    g = torch.Generator().manual_seed(1)
    ids = torch.randint(0, 40, (7, 5), generator=g)
    vals = torch.rand(7, 5, generator=g, dtype=torch.float64)
    _, y_v, _ = m.fm_terms(ids, vals)
    e = m.fm_v[ids] * vals[..., None]
    want = torch.zeros(7, dtype=torch.float64)
    for f in range(5):
        for h in range(f + 1, 5):
            want += (e[:, f] * e[:, h]).sum(-1)
    torch.testing.assert_close(y_v, want, atol=1e-12, rtol=1e-10)


@pytest.mark.parametrize("B,F,K,U", [(64, 39, 32, 500), (7, 5, 3, 11), (300, 39, 64, 4000), (1, 1, 16, 1)])
def test_native_cpu_embed_fm_matches_the_pytorch_expression(B, F, K, U, monkeypatch):
    """csrc/cpu/embed_fm_cpu.cc (what PS workers and the CPU eager back-end run) against the plain PyTorch expression
    of HVD.py:196-206: outputs, and gradients w.r.t. the distinct rows with an arbitrary downstream gradient."""
    import torch
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    from deepfm_b200.ops import embed_fm_cpu
    g = torch.Generator().manual_seed(B * 7 + K)
    inverse = torch.randint(0, U, (B, F), generator=g)
    if F > 1:
        inverse[0, 1] = inverse[0, 0]                     # a repeated row inside one example
    vals = torch.rand(B, F, generator=g)
    w0, v0 = torch.randn(U, generator=g), torch.randn(U, K, generator=g) * 0.3
    assert embed_fm_cpu.usable(w0, v0, inverse, vals)
    ge = torch.randn(B, F, K, generator=g)
    gy = torch.randn(2, B, generator=g)

    def run(native):
        w, v = w0.clone().requires_grad_(True), v0.clone().requires_grad_(True)
        if native:
            y_w, y_v, e = embed_fm_cpu.embed_fm(w, v, inverse, vals)
        else:
            y_w = (w[inverse] * vals).sum(1)
            e = v[inverse] * vals.unsqueeze(-1)
            y_v = 0.5 * (e.sum(1).pow(2) - e.pow(2).sum(1)).sum(1)
        ((y_w * gy[0]).sum() + (y_v * gy[1]).sum() + (e * ge).sum()).backward()
        return y_w.detach(), y_v.detach(), e.detach(), w.grad, v.grad

    for a, b in zip(run(True), run(False)):
        torch.testing.assert_close(a, b, atol=2e-5, rtol=1e-4)
    # only some outputs used downstream: the unused ones arrive as None in backward
    w, v = w0.clone().requires_grad_(True), v0.clone().requires_grad_(True)
    embed_fm_cpu.embed_fm(w, v, inverse, vals)[1].sum().backward()
    e = v0[inverse] * vals.unsqueeze(-1)
    want = torch.zeros_like(v0).index_add_(0, inverse.reshape(-1), ((e.sum(1, keepdim=True) - e) * vals.unsqueeze(-1)).reshape(-1, K))
    torch.testing.assert_close(v.grad, want, atol=2e-5, rtol=1e-4)
    assert w.grad is None or w.grad.abs().max() == 0
    # the model takes the native path on the host and the switch turns it off
    m = DeepFM(DeepFMConfig(F, max(U, 2), K, [8], [1.0]), seed=1)
    rows_w, rows_v = m.fm_w.detach()[:U].clone(), m.fm_v.detach()[:U].clone()
    a = m.forward_rows(rows_w, rows_v, inverse, vals)
    monkeypatch.setenv("DFM_CPU_NATIVE", "0")
    b = m.forward_rows(rows_w, rows_v, inverse, vals)
    torch.testing.assert_close(a, b, atol=1e-5, rtol=1e-5)


@pytest.mark.parametrize("rows,cols,keep,layer,step,seed", [(1024, 128, 0.5, 0, 1, 7), (300, 100, 0.9, 2, 12345, 2 ** 40 + 17),
                                                             (5, 3, 0.3, 1, 2, 0), (17, 64, 1.0, 0, 9, 3)])
def test_native_keep_mask_is_the_engines_philox_stream(rows, cols, keep, layer, step, seed, monkeypatch):
    """csrc/cpu dfcpu_keep_mask (what the PS workers draw their dropout from) is bit-equal to the tensor restatement of
    csrc/cuda/common.cuh dropout_keep_bits8 in utils/philox.py, which the GPU tests compare with the kernels."""
    from deepfm_b200.utils import philox
    a = philox.keep_mask(seed, step, layer, rows, cols, keep)
    monkeypatch.setenv("DFM_CPU_NATIVE", "0")
    b = philox.keep_mask(seed, step, layer, rows, cols, keep)
    assert a.dtype == torch.bool and a.shape == (rows, cols) and torch.equal(a, b)
    if keep < 1.0 and rows * cols > 1000:
        assert abs(a.float().mean().item() - keep) < 0.02
    if keep == 1.0:
        assert bool(a.all())


def test_serving_rejects_ids_outside_the_table(tmp_path):
    from deepfm_b200.utils.export import export_servable, load_servable
    m = DeepFM(DeepFMConfig(4, 50, 8, [8], [1.0]), seed=1)
    serve = load_servable(export_servable(m, str(tmp_path / "s")))
    ok = serve(np.array([[0, 1, 2, 49]]), np.ones((1, 4), np.float32))["prob"]
    assert ok.shape == (1,)
    for bad in (50, -1, 2 ** 40):
        with pytest.raises(IndexError, match="feature_size = 50"):
            serve(np.array([[0, 1, 2, bad]]), np.ones((1, 4), np.float32))


def test_native_adam_matches_the_tensor_formula(monkeypatch):
    """csrc/cpu dfcpu_adam_dense / dfcpu_adam_rows (what the parameter servers and the CPU trainer apply) against the
    eager TF formula in TFOptimizer, five steps, dense and lazy-row updates, vector and matrix variables."""
    g = torch.Generator().manual_seed(0)
    grads = [torch.randn(300, 32, generator=g) for _ in range(5)]
    rows = [torch.randperm(300, generator=g)[:40] for _ in range(5)]

    def run():
        opt = TFOptimizer("Adam", 0.01)
        w, v, d = torch.ones(300), torch.ones(300, 32), torch.ones(70000)
        for t in range(5):
            opt.apply_rows("fm_v", v, rows[t], grads[t][:40])
            opt.apply_rows("fm_w", w, rows[t], grads[t][:40, 0].contiguous())
            opt.apply_dense("d", d, grads[t].reshape(-1)[:9600].repeat(8)[:70000].contiguous())
            opt.finish_step()
        return w, v, d, opt.slots["fm_v/Adam"], opt.slots["fm_v/Adam_1"], opt.slots["d/Adam_1"]

    native = run()
    monkeypatch.setenv("DFM_CPU_NATIVE", "0")
    eager = run()
    for a, b in zip(native, eager):
        torch.testing.assert_close(a, b, atol=1e-6, rtol=1e-5)
    assert not torch.equal(native[1], torch.ones(300, 32))
    untouched = torch.ones(300, dtype=torch.bool)
    untouched[torch.cat(rows)] = False
    assert (native[1][untouched] == 1).all() and (native[3][untouched] == 0).all()


from hypothesis import given, settings, strategies as st  # noqa: E402


@settings(max_examples=60, deadline=None)
@given(seed=st.integers(0, 2 ** 31), rows=st.integers(1, 300), cols=st.integers(1, 40),
       V=st.sampled_from([7, 5000, 117581, 3_000_000]), skew=st.booleans())
def test_native_unique_inverse_is_torch_unique(seed, rows, cols, V, skew):
    """csrc/cpu/unique_cpu.cc (table pass + sort of the distinct ids only; radix sort above 4096 of them) returns exactly
    what torch.unique(return_inverse=True) returns, and leaves its table clean for the next call."""
    from deepfm_b200.ops.unique_cpu import unique_inverse
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, V, (rows, cols), generator=g)
    if skew:
        ids = (ids.double() / V).pow(6).mul(V).long().clamp_(0, V - 1)      # many repeats of small ids
    for _ in range(2):                                                       # second call: the scratch table was reset
        u, inv = unique_inverse(ids, V)
        tu, tinv = ids.unique(return_inverse=True)
        assert torch.equal(u, tu) and torch.equal(inv, tinv) and inv.shape == ids.shape
    with pytest.raises(IndexError, match="outside"):
        bad = ids.clone(); bad[-1, -1] = V
        unique_inverse(bad, V)
    u, inv = unique_inverse(ids, V)                                          # still clean after the error path
    assert torch.equal(u, ids.unique())


def test_native_unique_large_batch_uses_the_radix_path():
    from deepfm_b200.ops.unique_cpu import unique_inverse
    ids = torch.randint(0, 117581, (2048, 39), generator=torch.Generator().manual_seed(1))
    u, inv = unique_inverse(ids, 117581)
    tu, tinv = ids.unique(return_inverse=True)
    assert u.numel() > 4096 and torch.equal(u, tu) and torch.equal(inv, tinv)
