"""Fused engine (embed+FM, tcgen05 MLP fwd/bwd, wgrad, optimizers) vs the fp32 PyTorch oracle."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(B, layers, Ke=32, F=39, V=5000, opt="GD", lr=1.0, l2=0.0, mode="lazy", keep=None, seed=0):
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    from deepfm_b200.runtime.engine import FusedEngine
    torch.manual_seed(seed)
    cfg = DeepFMConfig(field_size=F, feature_size=V, embedding_size=Ke, deep_layers=layers,
                       dropout_keep=keep or [1.0] * len(layers), l2_reg=l2)
    m = DeepFM(cfg, seed=1).cuda()
    with torch.no_grad():
        m.fm_v.mul_(30); m.fm_w.mul_(30)
        for b in m.mlp_b: b.uniform_(-0.1, 0.1)
        m.out_b.fill_(0.05); m.fm_bias.fill_(-0.3)
    eng = FusedEngine(cfg, batch_max=B, optimizer=opt, lr=lr, table_update_mode=mode, seed=7)
    eng.load_from_module(m)
    g = torch.Generator(device="cuda").manual_seed(3)
    ids = torch.randint(1, V, (B, F), device="cuda", generator=g, dtype=torch.int64)
    ids[:, :13] = torch.arange(1, 14, device="cuda")
    ids[0, 14] = ids[0, 15]
    vals = torch.rand(B, F, device="cuda", generator=g)
    vals[:, 13:] = 1.0
    labels = (torch.rand(B, device="cuda", generator=g) < 0.3).float()
    return cfg, m, eng, ids, vals, labels


def _rt(x):
    """bf16 rounding with a straight-through gradient (what the tensor-core path sees)."""
    return x + (x.bfloat16().float() - x).detach()


def _emulated_logits(m, ids, vals):
    """The oracle with the engine's rounding points: bf16 MLP inputs/weights/activations,
    fp32 accumulation, fp32 FM terms and out layer."""
    y_w, y_v, e = m.fm_terms(ids, vals)
    h = _rt(e.reshape(e.shape[0], -1))
    for l in range(len(m.cfg.deep_layers)):
        h = _rt(torch.relu(h @ _rt(m.mlp_w[l]) + m.mlp_b[l]))
    return m.fm_bias + y_w + y_v + (h @ m.out_w).view(-1) + m.out_b


def _rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


@pytest.mark.parametrize("B,layers,Ke", [(256, [256, 128, 64], 32), (300, [128, 64, 32], 32), (2048, [256, 128, 64], 32),
                                         (384, [400, 400, 400], 64), (200, [64], 16), (256, [100, 50], 32)])
def test_forward_matches_oracle(B, layers, Ke):
    cfg, m, eng, ids, vals, labels = _setup(B, layers, Ke=Ke)
    prob = eng.forward(ids.int().contiguous(), vals, labels)
    with torch.no_grad():
        logits = m(ids, vals)
    torch.cuda.synchronize()
    assert _rel(eng.buffer("logit")[:B], logits) < 2e-2
    torch.testing.assert_close(prob, torch.sigmoid(logits), atol=1e-2, rtol=0)
    ref = torch.nn.functional.binary_cross_entropy_with_logits(logits, labels, reduction="sum").item()
    assert abs(eng.read_state()["loss_sum"] - ref) < 2e-2 * abs(ref) + 1e-3
    eng.close()


@pytest.mark.parametrize("B,layers,Ke,l2,mode", [(256, [256, 128, 64], 32, 0.0, "lazy"),
                                                 (300, [256, 128, 64], 32, 1e-3, "lazy"),
                                                 (2048, [256, 128, 64], 32, 0.0, "lazy"),
                                                 (384, [400, 400, 400], 64, 0.0, "lazy"),
                                                 (256, [128, 64, 32], 32, 1e-3, "dense_exact"),
                                                 (256, [100, 50], 32, 0.0, "lazy")])
def test_gd_step_equals_autograd_gradients(B, layers, Ke, l2, mode):
    cfg, m, eng, ids, vals, labels = _setup(B, layers, Ke=Ke, l2=l2, mode=mode)
    before = {k: v.clone() for k, v in eng.variables().items()}
    eng.train_step(ids.int().contiguous(), vals, labels)
    torch.cuda.synchronize()
    st = eng.read_state()
    assert st["step"] == 1
    m.loss(_emulated_logits(m, ids, vals), labels).backward()
    touched = torch.zeros(cfg.feature_size, dtype=torch.bool, device="cuda")
    touched[ids.unique()] = True
    for name, p in m.tf_variables().items():
        got = before[name] - eng.variables()[name]       # lr = 1
        ref = p.grad
        if name in ("fm_w", "fm_v") and mode == "lazy":
            ref = torch.where(touched if name == "fm_w" else touched[:, None], ref, torch.zeros_like(ref))
        assert _rel(got, ref) < 2e-2, name
    # bookkeeping is clean for the next step
    assert eng.buffer("G").abs().max().item() == 0
    assert eng.buffer("slot_of", dtype="i4").max().item() == -1
    eng.close()


@pytest.mark.parametrize("opt", ["Adam", "Adagrad", "Momentum", "ftrl", "GD"])
def test_optimizers_track_eager_reference(opt):
    """Three steps on the same batch: engine vs autograd + utils.optimizers (TF formulas)."""
    from deepfm_b200.utils.optimizers import TFOptimizer
    lr = 0.01
    cfg, m, eng, ids, vals, labels = _setup(256, [128, 64, 32], opt=opt, lr=lr, l2=1e-3)
    ref_opt = TFOptimizer(opt, lr)
    ids32 = ids.int().contiguous()
    uniq = ids.unique()
    for step in range(3):
        eng.train_step(ids32, vals, labels)
        m.zero_grad()
        m.data_loss(_emulated_logits(m, ids, vals), labels).backward()
        with torch.no_grad():
            for name, p in m.trainable_tf_variables().items():
                if name in ("fm_w", "fm_v"):       # lazy: touched rows, L2 on those rows
                    g = p.grad[uniq] + cfg.l2_reg * p[uniq]
                    ref_opt.apply_rows(name, p.data, uniq, g)
                else:
                    ref_opt.apply_dense(name, p.data, p.grad)
        ref_opt.finish_step()
    torch.cuda.synchronize()
    for name, p in m.trainable_tf_variables().items():
        got = eng.variables()[name]
        if name == "fm_v":
            got, p = got[uniq], p[uniq]
        assert _rel(got, p) < (8e-2 if opt in ("Adam", "Adagrad") else 3e-2), (opt, name)
    eng.close()


def test_dropout_masks_and_scaling():
    cfg, m, eng, ids, vals, labels = _setup(512, [256, 128, 64], keep=[0.5, 0.8, 1.0], opt="GD", lr=0.0)
    ids32 = ids.int().contiguous()
    eng.forward(ids32, vals, labels)
    pitch = (256 + 1 + 15) // 16 * 16
    h_eval = eng.buffer("act", 1, "bf16").view(-1, pitch)[:512, :256].float().clone()
    eng.train_step(ids32, vals, labels)
    h_train = eng.buffer("act", 1, "bf16").view(-1, pitch)[:512, :256].float().clone()
    torch.cuda.synchronize()
    pos = h_eval > 0
    kept = (h_train > 0) & pos
    frac = kept.sum().item() / pos.sum().item()
    assert 0.47 < frac < 0.53
    torch.testing.assert_close(h_train[kept], 2.0 * h_eval[kept], rtol=2e-2, atol=1e-3)
    eng.close()


def test_auc_histogram_matches_numpy():
    from deepfm_b200.runtime.engine import auc_histogram
    from deepfm_b200.utils.metrics import auc_from_hist, tf_auc_numpy
    g = torch.Generator(device="cuda").manual_seed(5)
    prob = torch.rand(10000, device="cuda", generator=g)
    prob[:5] = torch.tensor([0.0, 1.0, 0.5, 1.0 / 199, 198.0 / 199], device="cuda")
    label = (torch.rand(10000, device="cuda", generator=g) < prob * 0.7).float()
    hist = torch.zeros(2 * 201, dtype=torch.int64, device="cuda")
    auc_histogram(prob, label, hist)
    got = auc_from_hist(hist.cpu().numpy().reshape(2, 201))
    ref = tf_auc_numpy(label.cpu().numpy(), prob.cpu().numpy())
    assert abs(got - ref) < 1e-6


def test_pipelined_graph_step_matches_plain_launches():
    """Graph replay with the in-graph prefetch of the next batch == the same kernels launched one by one."""
    from deepfm_b200.data.loader import alloc_slab
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    from deepfm_b200.runtime.trainer import Trainer, TrainerOptions
    B, F, V = 256, 39, 5000
    cfg = DeepFMConfig(F, V, 32, (256, 128, 64), (0.5, 0.5, 0.5), l2_reg=1e-4)
    init = DeepFM(cfg, seed=11)
    g = torch.Generator().manual_seed(3)
    slabs = []
    for _ in range(7):
        slab, ids, vals, labels = alloc_slab(B, F, torch.int32, pin_memory=True)
        ids.copy_(torch.randint(1, V, (B, F), generator=g, dtype=torch.int32))
        vals.copy_(torch.rand(B, F, generator=g)); labels.copy_((torch.rand(B, generator=g) < 0.3).float())
        slabs.append(slab)
    dev = torch.device("cuda", 0)
    mk = lambda graph: Trainer(cfg, TrainerOptions(batch_size=B, optimizer="Adam", learning_rate=1e-3, backend="fused",
                                                   seed=5, cuda_graph=graph), device=dev, init_model=init)
    a, b = mk(True), mk(False)
    losses_a, losses_b = [], []
    for i, s in enumerate(slabs):
        nxt = slabs[i + 1] if i + 1 < len(slabs) and i != 3 else None      # step 4's batch is NOT prefetched
        a.train_step_host(s, nxt)
        losses_a.append(a.last_loss(wait_current=True))
        b.train_step_host(s)
        losses_b.append(b.last_loss(wait_current=True))
    # device-resident sources go through the second graph flavour
    d0, d1 = slabs[0].cuda(), slabs[1].cuda()
    a.train_step_device(d0, d1); a.train_step_device(d1, None)
    b.train_step_device(d0); b.train_step_device(d1)
    losses_a.append(a.last_loss(wait_current=True)); losses_b.append(b.last_loss(wait_current=True))
    torch.cuda.synchronize()
    # the per-CTA loss partials meet in a float atomicAdd: equal up to summation order
    np.testing.assert_allclose(losses_a, losses_b, rtol=1e-5)
    va, vb = a.variables(), b.variables()
    for k in va:
        # duplicate ids are accumulated with float atomics, so two runs agree only up to summation order
        torch.testing.assert_close(va[k], vb[k], rtol=1e-3, atol=2e-5, msg=k)
    a.close(); b.close()


def test_train_batches_pipelines_the_loader(tmp_path):
    """Trainer.train_batches (look-ahead + in-graph prefetch + deferred slot release) == one batch at a time."""
    from deepfm_b200.data.loader import BatchLoader
    from deepfm_b200.data.tfrecord import generate_synthetic
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    from deepfm_b200.runtime.trainer import Trainer, TrainerOptions
    path = str(tmp_path / "tr.tfrecords")
    generate_synthetic(path, 256 * 9 + 100, 39, 20000, n_dense=13, seed=3)
    cfg = DeepFMConfig(39, 20000, 32, (256, 128, 64), (1.0, 1.0, 1.0), l2_reg=1e-4)
    init = DeepFM(cfg, seed=2)
    dev = torch.device("cuda", 0)
    res = []
    for pipelined in (True, False):
        # plain SGD: Adam would turn summation-order noise on near-zero gradients into lr-sized differences
        tr = Trainer(cfg, TrainerOptions(batch_size=256, optimizer="GD", learning_rate=0.05, backend="fused", seed=1),
                     device=dev, init_model=init)
        loader = BatchLoader([path], 39, 256, num_epochs=2, drop_remainder=True, num_workers=2, num_slots=4)
        n = 0
        if pipelined:
            for _hb in tr.train_batches(loader):
                n += 1
        else:
            for hb in loader:
                tr.train_batch(hb)
                torch.cuda.synchronize()        # the slot is reused as soon as the next batch is requested
                n += 1
        loader.close()
        assert n == 18 and tr.global_step == 18
        res.append(({k: v.detach().clone() for k, v in tr.variables().items()}, tr.last_loss(wait_current=True)))
        tr.close()
    assert abs(res[0][1] - res[1][1]) < 1e-4
    for k in res[0][0]:
        torch.testing.assert_close(res[0][0][k], res[1][0][k], rtol=2e-3, atol=5e-5, msg=k)


def test_training_on_the_bundled_data_reaches_the_eager_auc(tmp_path, val_tfrecords):
    """SURVEY section 4, integration level: train on 80% of val.tfrecords, evaluate on the rest; the fused bf16 engine
    and the eager fp32 oracle must both learn (AUC well above 0.5) and agree."""
    import copy
    from deepfm_b200.data import tfrecord as T
    from deepfm_b200.data.loader import BatchLoader
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    from deepfm_b200.runtime.trainer import Trainer, TrainerOptions
    recs = list(T.iter_records(val_tfrecords))
    cut = int(len(recs) * 0.8)
    paths = {}
    for name, part in (("tr", recs[:cut]), ("va", recs[cut:])):
        paths[name] = str(tmp_path / f"{name}.tfrecords")
        with T.TFRecordWriter(paths[name]) as w:
            for r in part:
                w.write(r)
    cfg = DeepFMConfig(39, 117581, 32, (256, 128, 64), (0.5, 0.5, 0.5), l2_reg=1e-4)
    init = DeepFM(cfg, seed=1)
    aucs = {}
    for backend in ("eager", "fused"):
        # both arms use the engine's Philox dropout masks, so they differ by bf16 rounding and summation order only
        tr = Trainer(cfg, TrainerOptions(batch_size=256, optimizer="Adam", learning_rate=5e-4, backend=backend, seed=1,
                                         dropout_stream="engine"),
                     device=torch.device("cuda", 0), init_model=copy.deepcopy(init))
        for _ in range(3):
            loader = BatchLoader([paths["tr"]], 39, 256, num_epochs=1, drop_remainder=True, num_workers=2, num_slots=6)
            for _hb in tr.train_batches(loader):
                pass
            loader.close()
        loader = BatchLoader([paths["va"]], 39, 256, num_epochs=1, drop_remainder=False, num_workers=2)
        aucs[backend] = tr.evaluate(loader, auc_mode="both")
        loader.close()
        tr.close()
    assert aucs["eager"]["auc"] > 0.68 and aucs["fused"]["auc"] > 0.68, aucs
    # measured with shared masks: |diff| <= 0.0011 over three seeds (profiles/auc_parity_synthetic_r2.json)
    assert abs(aucs["eager"]["auc"] - aucs["fused"]["auc"]) < 0.005, aucs
    assert abs(aucs["fused"]["auc"] - aucs["fused"]["auc_exact"]) < 0.01, aucs


def test_out_of_range_ids_are_clamped_and_counted():
    """a feature id outside [0, feature_size) must not read or write outside the tables."""
    cfg, m, eng, ids, vals, labels = _setup(256, [128, 64, 32], V=5000)
    bad = ids.clone()
    bad[3, 20], bad[4, 21], bad[5, 22] = -5, 5000, 2 ** 31 - 1
    eng.train_step(bad.int().contiguous(), vals, labels)
    torch.cuda.synchronize()
    st = eng.read_state()
    assert st["bad_ids"] == 3
    ref = FusedEngineRef(cfg, m, bad.clamp(0, 4999), vals, labels)
    assert abs(st["loss_sum"] - ref) < 2e-2 * abs(ref) + 1e-3
    assert torch.isfinite(eng.variables()["fm_v"]).all()
    eng.close()


def FusedEngineRef(cfg, m, ids, vals, labels):
    with torch.no_grad():
        return torch.nn.functional.binary_cross_entropy_with_logits(m(ids, vals), labels, reduction="sum").item()


# ---------------------------------------------------------------------------------------------
# batch norm (FC -> ReLU -> BN -> dropout, HVD.py:224-235 + 306-328) and square loss on the fused path
# ---------------------------------------------------------------------------------------------
def _setup_bn(B, layers, loss_type="log_loss", V=5000, Ke=32, F=39, lr=1.0):
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    from deepfm_b200.runtime.engine import FusedEngine
    cfg = DeepFMConfig(field_size=F, feature_size=V, embedding_size=Ke, deep_layers=layers,
                       dropout_keep=[1.0] * len(layers), l2_reg=0.0, batch_norm=True, batch_norm_decay=0.9,
                       loss_type=loss_type)
    m = DeepFM(cfg, seed=1).cuda()
    g = torch.Generator(device="cuda").manual_seed(9)
    with torch.no_grad():
        m.fm_v.mul_(30); m.fm_w.mul_(30)
        for l in range(len(layers)):
            m.mlp_b[l].uniform_(-0.1, 0.1)
            m.bn_gamma[l].copy_(0.5 + torch.rand(layers[l], device="cuda", generator=g))
            m.bn_beta[l].copy_(0.2 * torch.randn(layers[l], device="cuda", generator=g))
            getattr(m, f"bn_mean_{l}").copy_(0.1 * torch.rand(layers[l], device="cuda", generator=g))
            getattr(m, f"bn_var_{l}").copy_(0.5 + torch.rand(layers[l], device="cuda", generator=g))
        m.out_b.fill_(0.05); m.fm_bias.fill_(-0.3)
    eng = FusedEngine(cfg, batch_max=B, optimizer="GD", lr=lr, seed=7)
    eng.load_from_module(m)
    ids = torch.randint(1, V, (B, F), device="cuda", generator=g, dtype=torch.int64)
    ids[:, :13] = torch.arange(1, 14, device="cuda")
    vals = torch.rand(B, F, device="cuda", generator=g)
    vals[:, 13:] = 1.0
    labels = (torch.rand(B, device="cuda", generator=g) < 0.3).float()
    return cfg, m, eng, ids, vals, labels


def _emulated_logits_bn(m, ids, vals, training):
    """Oracle with the engine's rounding points: bf16 a = relu(.), fp32 statistics of the rounded a, bf16 BN output."""
    from deepfm_b200.models.deepfm import BN_EPS
    y_w, y_v, e = m.fm_terms(ids, vals)
    h = _rt(e.reshape(e.shape[0], -1))
    for l in range(len(m.cfg.deep_layers)):
        a = _rt(torch.relu(h @ _rt(m.mlp_w[l]) + m.mlp_b[l]))
        if training:
            mean, var = a.mean(0), a.var(0, unbiased=False)
        else:
            mean, var = getattr(m, f"bn_mean_{l}"), getattr(m, f"bn_var_{l}")
        h = _rt((a - mean) * torch.rsqrt(var + BN_EPS) * m.bn_gamma[l] + m.bn_beta[l])
    return m.fm_bias + y_w + y_v + (h @ m.out_w).view(-1) + m.out_b


@pytest.mark.parametrize("B,layers", [(256, [256, 128, 64]), (300, [100, 50])])
def test_batch_norm_inference_uses_moving_statistics(B, layers):
    cfg, m, eng, ids, vals, labels = _setup_bn(B, layers)
    prob = eng.forward(ids.int().contiguous(), vals, labels)
    with torch.no_grad():
        logits = m(ids, vals, training=False)
    torch.cuda.synchronize()
    assert _rel(eng.buffer("logit")[:B], logits) < 2e-2
    torch.testing.assert_close(prob, torch.sigmoid(logits), atol=1e-2, rtol=0)
    eng.close()


@pytest.mark.parametrize("B,layers,loss_type", [(256, [256, 128, 64], "log_loss"), (300, [100, 50], "log_loss"),
                                                (256, [128, 64], "square_loss")])
def test_batch_norm_training_step_equals_autograd(B, layers, loss_type):
    cfg, m, eng, ids, vals, labels = _setup_bn(B, layers, loss_type)
    before = {k: v.clone() for k, v in eng.variables().items()}
    eng.train_step(ids.int().contiguous(), vals, labels)
    torch.cuda.synchronize()
    logits = _emulated_logits_bn(m, ids, vals, training=True)
    m.loss(logits, labels).backward()
    loss_ref = (m.data_loss(logits, labels) * B).item()
    assert abs(eng.read_state()["loss_sum"] - loss_ref) < 2e-2 * abs(loss_ref) + 1e-3
    touched = torch.zeros(cfg.feature_size, dtype=torch.bool, device="cuda")
    touched[ids.unique()] = True
    for name, p in m.trainable_tf_variables().items():
        got = before[name] - eng.variables()[name]       # GD, lr = 1
        ref = p.grad
        if name in ("fm_w", "fm_v"):
            ref = torch.where(touched if name == "fm_w" else touched[:, None], ref, torch.zeros_like(ref))
        assert _rel(got, ref) < 3e-2, name
    # moving averages moved towards the batch statistics (decay 0.9), like contrib batch_norm with updates_collections=None
    with torch.no_grad():
        m2 = type(m)(cfg, seed=1).cuda()
        m2.load_tf_variables({k: v for k, v in before.items()})
        m2(ids, vals, training=True)
    for l in range(len(layers)):
        for stat in ("moving_mean", "moving_variance"):
            k = f"Deep-part/bn_{l}/{stat}"
            torch.testing.assert_close(eng.variables()[k], m2.tf_variables()[k], rtol=3e-2, atol=3e-3)
    eng.close()


def test_square_loss_without_batch_norm():
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    from deepfm_b200.runtime.engine import FusedEngine
    B, V, layers = 256, 5000, [128, 64, 32]
    cfg = DeepFMConfig(39, V, 32, layers, [1.0] * 3, l2_reg=0.0, loss_type="square_loss")
    m = DeepFM(cfg, seed=1).cuda()
    with torch.no_grad():
        m.fm_v.mul_(30); m.fm_w.mul_(30)
    eng = FusedEngine(cfg, batch_max=B, optimizer="GD", lr=1.0, seed=7)
    eng.load_from_module(m)
    g = torch.Generator(device="cuda").manual_seed(3)
    ids = torch.randint(1, V, (B, 39), device="cuda", generator=g, dtype=torch.int64)
    vals = torch.rand(B, 39, device="cuda", generator=g)
    labels = (torch.rand(B, device="cuda", generator=g) < 0.3).float()
    before = {k: v.clone() for k, v in eng.variables().items()}
    eng.train_step(ids.int().contiguous(), vals, labels)
    torch.cuda.synchronize()
    logits = _emulated_logits(m, ids, vals)
    m.loss(logits, labels).backward()
    ref_loss = ((torch.sigmoid(logits) - labels) ** 2).sum().item()
    assert abs(eng.read_state()["loss_sum"] - ref_loss) < 2e-2 * ref_loss + 1e-3
    for name in ("Deep-part/mlp0/weights", "Deep-part/deep_out/weights", "fm_bias"):
        assert _rel(before[name] - eng.variables()[name], dict(m.tf_variables())[name].grad) < 2e-2, name
    eng.close()


def test_python_philox_reproduces_the_engine_dropout_masks():
    """utils/philox.keep_mask == the keep decisions the layer kernels make (same seed, step, layer, row, column)."""
    from deepfm_b200.utils.philox import keep_mask
    cfg, m, eng, ids, vals, labels = _setup(300, [256, 128, 64], keep=[0.5, 0.8, 0.3], opt="GD", lr=0.0)
    ids32 = ids.int().contiguous()
    eng.forward(ids32, vals, labels)
    evals = [eng.buffer("act", l + 1, "bf16").clone() for l in range(3)]
    eng.train_step(ids32, vals, labels)                      # optimizer step 1, seed 7 (see _setup)
    torch.cuda.synchronize()
    for l, (n, keep) in enumerate(zip([256, 128, 64], [0.5, 0.8, 0.3])):
        pitch = (n + 1 + 15) // 16 * 16
        h = eng.buffer("act", l + 1, "bf16").view(-1, pitch)[:300, :n].float()
        mask = keep_mask(7, 1, l, 300, n, keep, device="cuda")
        assert abs(mask.float().mean().item() - keep) < 0.02
        assert (h[~mask] == 0).all()                          # dropped where the mirror says so
        if l == 0:                                            # layer 0's input does not depend on earlier masks
            he = evals[0].view(-1, pitch)[:300, :n].float()
            live = he > 0
            assert ((h > 0) == (live & mask)).all()
    eng.close()


def test_gradients_elementwise_and_against_the_unrounded_oracle():
    """a Frobenius-norm check would let a few wrong rows of a 100k-row table through.  Here every touched
    table row and every dense weight is compared element by element with the rounding-emulating oracle, and the whole
    update is also compared with the plain fp32 oracle under a stated bf16 bound."""
    B, V = 512, 20000
    cfg, m, eng, ids, vals, labels = _setup(B, [256, 128, 64], V=V, l2=1e-3)
    before = {k: v.clone() for k, v in eng.variables().items()}
    eng.train_step(ids.int().contiguous(), vals, labels)
    torch.cuda.synchronize()
    got = {k: before[k] - eng.variables()[k] for k in before}          # GD, lr = 1: the gradient
    uniq = ids.unique()
    l2 = cfg.l2_reg
    # (a) oracle with the engine's rounding points
    m.zero_grad()
    m.data_loss(_emulated_logits(m, ids, vals), labels).backward()
    for name, p in m.tf_variables().items():
        ref = p.grad.clone()
        if name in ("fm_w", "fm_v"):
            ref[uniq] += l2 * p.detach()[uniq]
            g, r = got[name][uniq], ref[uniq]
            untouched = torch.ones(V, dtype=torch.bool, device="cuda"); untouched[uniq] = False
            assert got[name][untouched].abs().max().item() == 0, name      # lazy: rows not in the batch do not move
        else:
            g, r = got[name], ref
        scale = r.abs().max().item()
        torch.testing.assert_close(g, r, rtol=5e-2, atol=2e-2 * scale, msg=name)
    # (b) the plain fp32 oracle.  bf16 inputs, weights, activations and dz (8 mantissa bits each, chained through three
    #     ReLU layers whose units near zero can flip) cost a few percent of a gradient tensor's norm: measured here
    #     0.067 for Deep-part/mlp0/weights (the longest chain), <= 0.03 for the others; the bound is 0.1
    m.zero_grad()
    m.data_loss(m(ids, vals, training=False), labels).backward()
    for name, p in m.tf_variables().items():
        ref = p.grad.clone()
        if name in ("fm_w", "fm_v"):
            ref[uniq] += l2 * p.detach()[uniq]
            ref_sel, got_sel = ref[uniq], got[name][uniq]
        else:
            ref_sel, got_sel = ref, got[name]
        assert _rel(got_sel, ref_sel) < 1e-1, name
    eng.close()


def test_partial_batch_between_graph_replays():
    """Trainer.train_batch with fewer rows than batch_size goes through plain launches, which rewrite the
    device-side row count the captured graph reads.  The count must be back to the full batch for the next replay, and
    the partial step's loss is averaged over its own rows."""
    from deepfm_b200.data.loader import HostBatch, alloc_slab
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    from deepfm_b200.runtime.trainer import Trainer, TrainerOptions
    B, F, V = 256, 39, 5000
    cfg = DeepFMConfig(F, V, 32, (128, 64), (1.0, 1.0), l2_reg=0.0)
    g = torch.Generator().manual_seed(3)
    slab, ids, vals, labels = alloc_slab(B, F, torch.int32, pin_memory=True)
    ids.copy_(torch.randint(1, V, (B, F), generator=g, dtype=torch.int32))
    vals.copy_(torch.rand(B, F, generator=g)); labels.copy_((torch.rand(B, generator=g) < 0.3).float())
    tr = Trainer(cfg, TrainerOptions(batch_size=B, optimizer="GD", learning_rate=0.0, backend="fused", seed=5,
                                     cuda_graph=True), device=torch.device("cuda", 0), init_model=DeepFM(cfg, seed=11))
    tr.train_step_host(slab)
    full = tr.last_loss(wait_current=True)
    n = 100
    tr.train_batch(HostBatch(ids[:n].clone(), vals[:n].clone(), labels[:n].clone(), n))
    part = tr.last_loss(wait_current=True)
    assert tr.engine.read_state()["count"] == B                      # restored for the graph
    m = tr.export_model()
    with torch.no_grad():
        ref_part = float(m.data_loss(m(ids[:n].long(), vals[:n]), labels[:n]))
    assert abs(part - ref_part) < 2e-2 * abs(ref_part) + 1e-3        # mean over the 100 rows, not over 256
    tr.train_step_host(slab)                                         # weights are frozen (lr 0): same batch, same loss
    again = tr.last_loss(wait_current=True)
    assert abs(again - full) < 1e-5 * abs(full) + 1e-6
    tr.close()


def test_servable_answered_by_the_fused_kernels(tmp_path):
    """The export of HVD.py:477-493 served through the sm_100a inference kernels (load_servable(backend="fused")):
    same signature and the same probabilities as the eager loader, request larger than one engine batch."""
    import numpy as np
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    from deepfm_b200.utils.export import export_servable, load_servable
    cfg = DeepFMConfig(field_size=39, feature_size=5000, embedding_size=32, deep_layers=[128, 64, 32],
                       dropout_keep=[0.5] * 3)
    m = DeepFM(cfg, seed=3)
    with torch.no_grad():
        m.fm_v.mul_(30); m.fm_w.mul_(30); m.fm_bias.fill_(-0.2)
    path = export_servable(m, str(tmp_path / "export"))
    eager = load_servable(path)
    fused = load_servable(str(tmp_path / "export"), device="cuda", backend="fused", max_batch=256)
    rng = np.random.default_rng(0)
    ids = rng.integers(1, 5000, (700, 39))                  # int64, like the serving signature
    vals = rng.random((700, 39)).astype(np.float32)
    a, b = eager(ids, vals)["prob"], fused(ids, vals)["prob"]
    assert b.shape == (700,) and b.dtype == np.float32
    assert np.abs(a - b).max() < 1e-2 and a.std() > 0.01
    fused.engine.close()
