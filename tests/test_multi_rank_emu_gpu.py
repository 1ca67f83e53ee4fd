"""The multi-rank protocol of csrc/cuda/comm.cu on ONE GPU: N engines act as the ranks of a synchronous data-parallel
job (hvd.DistributedOptimizer, HVD.py:296) and the two collective kernels -- dense all-reduce + optimizer, sparse
owner-sharded exchange + lazy update -- run once, cooperatively, over all virtual ranks (runtime.engine.EmulatedRanks).
Oracle: plain PyTorch fp32 autograd per rank, gradients averaged over ranks, TF-formula optimizers."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rt(x):
    return x + (x.bfloat16().float() - x).detach()


def _emulated_logits(m, ids, vals):
    y_w, y_v, e = m.fm_terms(ids, vals)
    h = _rt(e.reshape(e.shape[0], -1))
    for l in range(len(m.cfg.deep_layers)):
        h = _rt(torch.relu(h @ _rt(m.mlp_w[l]) + m.mlp_b[l]))
    return m.fm_bias + y_w + y_v + (h @ m.out_w).view(-1) + m.out_b


def _rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


def _make(n, B, layers, Ke, V, opt, lr, l2, mode="lazy", F=39):
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    from deepfm_b200.runtime.engine import EmulatedRanks
    cfg = DeepFMConfig(field_size=F, feature_size=V, embedding_size=Ke, deep_layers=layers,
                       dropout_keep=[1.0] * len(layers), l2_reg=l2)
    m = DeepFM(cfg, seed=1).cuda()
    with torch.no_grad():
        m.fm_v.mul_(30); m.fm_w.mul_(30)
        for b in m.mlp_b: b.uniform_(-0.1, 0.1)
        m.out_b.fill_(0.05); m.fm_bias.fill_(-0.3)
    emu = EmulatedRanks(cfg, n, B, optimizer=opt, lr=lr, table_update_mode=mode, seed=7)
    emu.load_from_module(m)
    return cfg, m, emu


def _batches(n, B, F, V, seed):
    out = []
    for r in range(n):
        g = torch.Generator(device="cuda").manual_seed(seed * 100 + r)
        ids = torch.randint(1, V, (B, F), device="cuda", generator=g, dtype=torch.int64)
        ids[:, :13] = torch.arange(1, 14, device="cuda")          # the Criteo numeric fields: every rank, every row
        ids[0, 14] = ids[0, 15]
        vals = torch.rand(B, F, device="cuda", generator=g)
        vals[:, 13:] = 1.0
        labels = (torch.rand(B, device="cuda", generator=g) < 0.3).float()
        out.append((ids, vals, labels))
    return out


def _oracle_grads(m, batches):
    """Dense gradients averaged over ranks; table gradients averaged over ranks (dense [V, ...] tensors)."""
    n = len(batches)
    acc = {}
    for ids, vals, labels in batches:
        m.zero_grad()
        m.data_loss(_emulated_logits(m, ids, vals), labels).backward()
        for name, p in m.trainable_tf_variables().items():
            acc[name] = acc.get(name, 0) + p.grad.detach().clone() / n
    return acc


def _assert_replicas_identical(emu):
    e0 = emu.engines[0]
    for e in emu.engines[1:]:
        for k, v in e0.variables().items():
            assert torch.equal(v, e.variables()[k]), f"replica {e.rank} differs in {k}"


@pytest.mark.parametrize("n,B,layers,Ke,V", [(2, 256, [256, 128, 64], 32, 5000), (4, 200, [128, 64, 32], 32, 3001),
                                             (8, 128, [128, 64], 32, 20011), (2, 256, [400, 400], 64, 5000),
                                             (4, 128, [64, 32], 16, 4099), (3, 128, [64, 32], 32, 4099)])
def test_gd_step_equals_rank_averaged_autograd(n, B, layers, Ke, V):
    l2 = 1e-3
    cfg, m, emu = _make(n, B, layers, Ke, V, "GD", 1.0, l2)
    batches = _batches(n, B, cfg.field_size, V, seed=3)
    before = {k: v.clone() for k, v in emu.engines[0].variables().items()}
    emu.train_step([(i.int().contiguous(), v, l) for i, v, l in batches])
    torch.cuda.synchronize()
    _assert_replicas_identical(emu)
    grads = _oracle_grads(m, batches)
    touched = torch.zeros(V, dtype=torch.bool, device="cuda")
    for ids, _, _ in batches:
        touched[ids.unique()] = True
    for name, p in m.tf_variables().items():
        got = before[name] - emu.engines[0].variables()[name]             # lr = 1
        ref = grads[name]
        if name in ("fm_w", "fm_v"):                                       # lazy: touched rows, L2 on those rows
            mask = touched if name == "fm_w" else touched[:, None]
            ref = torch.where(mask, ref + l2 * p.detach(), torch.zeros_like(ref))
            torch.testing.assert_close(got[~touched], torch.zeros_like(got[~touched]), atol=0, rtol=0)
        assert _rel(got, ref) < 2e-2, name
    for e in emu.engines:                                                  # bookkeeping is clean for the next step
        assert e.buffer("G").abs().max().item() == 0
        assert e.buffer("slot_of", dtype="i4").max().item() == -1
    emu.close()


@pytest.mark.parametrize("n,opt", [(2, "Adam"), (4, "Adam"), (8, "Adagrad"), (2, "Momentum"), (2, "ftrl")])
def test_optimizers_over_three_steps(n, opt):
    """Slots of a table row live on its owner (id % n); parameters everywhere."""
    from deepfm_b200.utils.optimizers import TFOptimizer
    B, V, lr, l2 = 128, 4001, 0.01, 1e-3
    cfg, m, emu = _make(n, B, [128, 64, 32], 32, V, opt, lr, l2)
    ref_opt = TFOptimizer(opt, lr)
    for step in range(3):
        batches = _batches(n, B, cfg.field_size, V, seed=10 + step)
        emu.train_step([(i.int().contiguous(), v, l) for i, v, l in batches])
        grads = _oracle_grads(m, batches)
        uniq = torch.cat([b[0].reshape(-1) for b in batches]).unique()
        with torch.no_grad():
            for name, p in m.trainable_tf_variables().items():
                if name in ("fm_w", "fm_v"):
                    ref_opt.apply_rows(name, p.data, uniq, grads[name][uniq] + l2 * p[uniq])
                else:
                    ref_opt.apply_dense(name, p.data, grads[name])
        ref_opt.finish_step()
    torch.cuda.synchronize()
    _assert_replicas_identical(emu)
    tol = 8e-2 if opt in ("Adam", "Adagrad") else 3e-2
    for name, p in m.trainable_tf_variables().items():
        assert _rel(emu.engines[0].variables()[name], p) < tol, (opt, name)
    emu.close()


def test_dense_exact_two_ranks():
    """Reference math (l2 over the whole table, every row updated every step; HVD.py:266-267) on the multi-rank path."""
    n, B, V, l2 = 2, 128, 3001, 1e-3
    cfg, m, emu = _make(n, B, [128, 64, 32], 32, V, "GD", 1.0, l2, mode="dense_exact")
    batches = _batches(n, B, cfg.field_size, V, seed=5)
    before = {k: v.clone() for k, v in emu.engines[0].variables().items()}
    emu.train_step([(i.int().contiguous(), v, l) for i, v, l in batches])
    torch.cuda.synchronize()
    _assert_replicas_identical(emu)
    grads = _oracle_grads(m, batches)
    for name, p in m.tf_variables().items():
        got = before[name] - emu.engines[0].variables()[name]
        ref = grads[name] + (l2 * p.detach() if name in ("fm_w", "fm_v") else 0)
        assert _rel(got, ref) < 2e-2, name
    assert emu.engines[0].grad_table.abs().max().item() == 0
    emu.close()


def test_many_steps_keep_replicas_identical():
    """200 back-to-back steps (flag epochs advance, inboxes are reused): replicas never drift, loss falls."""
    n, B, V = 4, 256, 10007
    cfg, m, emu = _make(n, B, [256, 128, 64], 32, V, "Adam", 2e-3, 1e-4)
    first = last = None
    for step in range(200):
        batches = _batches(n, B, cfg.field_size, V, seed=step % 5)
        emu.train_step([(i.int().contiguous(), v, l) for i, v, l in batches])
        if step in (0, 199):
            torch.cuda.synchronize()
            loss = sum(e.read_state()["loss_sum"] for e in emu.engines) / (n * B)
            first, last = (loss, last) if step == 0 else (first, loss)
    torch.cuda.synchronize()
    _assert_replicas_identical(emu)
    assert last < first
    emu.close()


def test_watchdog_trips_when_a_rank_is_out_of_step():
    """A rank that waits for a peer's flag longer than DFM_COMM_TIMEOUT_S prints what it waits for and traps -- in a
    subprocess, because a trap leaves the CUDA context unusable."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, os.path.join(root, "scripts", "gpu_watchdog_trip.py")], capture_output=True,
                       text=True, timeout=180)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "[dfm comm watchdog]" in p.stdout + p.stderr and "watchdog fired" in p.stdout
