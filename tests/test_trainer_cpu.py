"""Trainer on the eager back-end: learning, checkpoint/resume, CLI end-to-end, sync-DP invariants (gloo)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from deepfm_b200.data import tfrecord as T
from deepfm_b200.data.loader import BatchLoader, HostBatch
from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.runtime.trainer import Trainer, TrainerOptions

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _batches(n, B, F=8, V=200, seed=0):
    rng = np.random.default_rng(seed)
    w = rng.standard_normal(V) * 1.5
    out = []
    for _ in range(n):
        ids = rng.integers(1, V, (B, F))
        vals = np.ones((B, F), np.float32)
        y = (rng.random(B) < 1 / (1 + np.exp(-w[ids].sum(1) * 0.5))).astype(np.float32)
        out.append(HostBatch(torch.from_numpy(ids).int(), torch.from_numpy(vals), torch.from_numpy(y), B))
    return out


def test_eager_trainer_learns_and_evaluates():
    cfg = DeepFMConfig(8, 200, 8, [32, 16], [1.0, 1.0], l2_reg=1e-5)
    tr = Trainer(cfg, TrainerOptions(batch_size=128, optimizer="Adam", learning_rate=0.01, backend="eager", seed=1),
                 device="cpu")
    data = _batches(60, 128)
    before = tr.evaluate(data[50:], auc_mode="both")
    for _ in range(4):
        for hb in data[:50]:
            tr.train_batch(hb)
    after = tr.evaluate(data[50:], auc_mode="both")
    assert tr.global_step == 200 and after["examples"] == 1280
    assert after["data_loss"] < before["data_loss"] - 0.02 and after["auc"] > 0.62 and after["auc"] > before["auc"]
    assert abs(after["auc"] - after["auc_exact"]) < 5e-3
    probs = np.concatenate(list(tr.predict(data[50:52])))
    assert probs.shape == (256,) and (probs > 0).all() and (probs < 1).all()


@pytest.mark.parametrize("mode", ["lazy", "dense_exact"])
def test_table_update_modes(mode):
    cfg = DeepFMConfig(8, 200, 8, [16], [1.0], l2_reg=0.01)
    tr = Trainer(cfg, TrainerOptions(batch_size=32, optimizer="GD", learning_rate=0.1, table_update_mode=mode,
                                     backend="eager", seed=1), device="cpu")
    hb = _batches(1, 32)[0]
    before = tr.model.fm_v.detach().clone()
    tr.train_batch(hb)
    moved = (tr.model.fm_v.detach() != before).any(1)
    touched = torch.zeros(200, dtype=torch.bool)
    touched[hb.feat_ids.long().unique()] = True
    if mode == "lazy":
        assert torch.equal(moved, touched)                 # untouched rows are not even decayed
    else:
        assert moved.all()                                 # l2 * V reaches every row (reference math)
        untouched = ~touched
        torch.testing.assert_close(tr.model.fm_v.detach()[untouched], before[untouched] * (1 - 0.1 * 0.01))


def test_checkpoint_resume_continues_global_step(tmp_path):
    from deepfm_b200.utils.checkpoint import CheckpointManager
    cfg = DeepFMConfig(8, 200, 8, [16], [1.0])
    opts = TrainerOptions(batch_size=64, optimizer="Adam", learning_rate=0.01, backend="eager", seed=1)
    data = _batches(6, 64)
    a = Trainer(cfg, opts, device="cpu")
    for hb in data[:3]:
        a.train_batch(hb)
    mgr = CheckpointManager(str(tmp_path), keep_max=5)
    mgr.save(a.global_step, a.checkpoint_tensors())
    for hb in data[3:]:
        a.train_batch(hb)
    b = Trainer(cfg, TrainerOptions(batch_size=64, optimizer="Adam", learning_rate=0.01, backend="eager", seed=99),
                device="cpu")
    step, tensors = mgr.restore_latest()
    assert step == 3 and "fm_v/Adam" in tensors and "Deep-part/mlp0/weights/Adam_1" in tensors
    assert abs(float(tensors["beta1_power"]) - 0.9 ** 4) < 1e-7
    b.load_checkpoint_tensors(step, tensors)
    for hb in data[3:]:
        b.train_batch(hb)
    assert b.global_step == 6
    for k, v in a.variables().items():
        torch.testing.assert_close(b.variables()[k], v, rtol=1e-6, atol=1e-7)


def test_cli_train_eval_infer_export(tmp_path, val_tfrecords):
    recs = list(T.iter_records(val_tfrecords))[:3000]
    (tmp_path / "train").mkdir()
    (tmp_path / "val").mkdir()
    for name, part in (("train/tr_0.tfrecords", recs[:2000]), ("val/va_0.tfrecords", recs[2000:]),
                       ("val/te_0.tfrecords", recs[2000:2300])):
        with T.TFRecordWriter(str(tmp_path / name)) as w:
            for r in part:
                w.write(r)
    env = dict(os.environ, PYTHONPATH=ROOT)
    for k in ("SM_HOSTS", "SM_CURRENT_HOST", "SM_CHANNELS", "RANK", "WORLD_SIZE"):
        env.pop(k, None)
    base = [sys.executable, "-m", "deepfm_b200.train", "--field_size", "39", "--feature_size", "117581",
            "--deep_layers", "32,16", "--dropout", "0.8,0.8", "--embedding_size", "8", "--batch_size", "200",
            "--training_data_dir", str(tmp_path / "train"), "--val_data_dir", str(tmp_path / "val"),
            "--checkpoint_dir", str(tmp_path / "ckpt"), "--servable_model_dir", str(tmp_path / "export"),
            "--backend", "eager", "--perform_shuffle", "False", "--log_steps", "5"]
    r = subprocess.run(base + ["--task_type", "train", "--num_epochs", "2"], env=env, capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ignored unknown arguments: --perform_shuffle False" in r.stdout
    assert "epoch 2" in r.stdout and "'auc'" in r.stdout and "model.ckpt-20" in r.stdout
    assert os.path.exists(tmp_path / "ckpt" / "checkpoint") and os.listdir(tmp_path / "export")
    r = subprocess.run(base + ["--task_type", "eval"], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "Restoring parameters" in r.stdout and "'examples': 1000" in r.stdout, r.stdout + r.stderr
    r = subprocess.run(base + ["--task_type", "infer"], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    preds = [float(x) for x in open(tmp_path / "val" / "pred.txt")]
    assert len(preds) == 300 and all(0 < p < 1 for p in preds)


# ---------------------------------------------------------------- synchronous data parallel (gloo)
def _dp_worker(rank, world, port, out_dir, opt):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = DeepFMConfig(8, 200, 8, [16, 8], [1.0, 1.0], l2_reg=1e-3)
    tr = Trainer(cfg, TrainerOptions(batch_size=32, optimizer=opt, learning_rate=0.02, backend="eager", seed=5),
                 device="cpu")
    data = _batches(4, 32 * world, seed=3)
    for hb in data:                                  # rank r takes rows r::world of the global batch
        sl = slice(rank, None, world)
        tr.train_batch(HostBatch(hb.feat_ids[sl].contiguous(), hb.feat_vals[sl].contiguous(),
                                 hb.labels[sl].contiguous(), 32))
    torch.save({k: v.detach().clone() for k, v in tr.variables().items()}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("opt", ["GD", "Adam"])
def test_sync_dp_equals_big_batch_with_scaled_lr(tmp_path, opt):
    """N ranks x batch B with lr*N (HVD.py:171) == 1 rank x batch N*B with the same effective lr."""
    world = 2
    mp.spawn(_dp_worker, args=(world, _free_port(), str(tmp_path), opt), nprocs=world, join=True)
    r0, r1 = (torch.load(tmp_path / f"r{i}.pt") for i in range(world))
    for k in r0:                                     # replicas never drift
        assert torch.equal(r0[k], r1[k]), k
    cfg = DeepFMConfig(8, 200, 8, [16, 8], [1.0, 1.0], l2_reg=1e-3)
    single = Trainer(cfg, TrainerOptions(batch_size=64, optimizer=opt, learning_rate=0.02 * world, backend="eager",
                                         seed=5), device="cpu")
    for hb in _batches(4, 64, seed=3):
        single.train_batch(hb)
    for k, v in single.variables().items():
        torch.testing.assert_close(r0[k], v.detach(), rtol=2e-4, atol=2e-6)


def test_train_batches_consumes_every_batch_once_in_order(tmp_path):
    """Regression: without the pipelined path a look-ahead fetch used to hand the current batch's ring slot back to the
    parser threads before it was trained on (batches skipped / seen twice).  With frozen weights the per-step loss is
    a fingerprint of the batch: it must equal the loss of the file's batches in file order."""
    import numpy as np
    import torch
    from deepfm_b200.data import tfrecord as T
    from deepfm_b200.data.loader import BatchLoader, HostBatch
    from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
    from deepfm_b200.runtime.trainer import Trainer, TrainerOptions
    path = str(tmp_path / "tr.tfrecords")
    B, n = 32, 32 * 40
    T.generate_synthetic(path, n, 39, 5000, n_dense=13, seed=7)
    cfg = DeepFMConfig(39, 5000, 8, (16, 8), (1.0, 1.0), l2_reg=0.0)
    init = DeepFM(cfg, seed=1)
    with torch.no_grad():
        init.fm_v.mul_(50); init.fm_w.mul_(50)
    tr = Trainer(cfg, TrainerOptions(batch_size=B, optimizer="GD", learning_rate=0.0, backend="eager", seed=1),
                 device="cpu", init_model=init)
    loader = BatchLoader([path], 39, B, num_epochs=1, drop_remainder=True, num_workers=3, num_slots=3)
    got = []
    for _ in tr.train_batches(loader):
        got.append(tr.last_loss(wait_current=True))
    loader.close()
    recs = [T.parse_example(r, 39) for r in T.iter_records(path)]
    want = []
    for b in range(n // B):
        part = recs[b * B:(b + 1) * B]
        ids = torch.from_numpy(np.stack([p[1] for p in part])); vals = torch.from_numpy(np.stack([p[2] for p in part]))
        labels = torch.tensor([p[0] for p in part])
        with torch.no_grad():
            want.append(float(init.data_loss(init(ids, vals), labels)))
    np.testing.assert_allclose(got, want, rtol=1e-5)


def _gather_proc(rank, world, port, out_dir):
    import os
    import types
    import torch
    import torch.distributed as dist
    from deepfm_b200.parallel.sync_dp import gather_owner_sharded_slots
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    V, K = 23, 4
    # every rank holds current values only in the rows it owns (id % world == rank); the rest is stale (-1)
    def shard(shape_tail):
        t = torch.full((V,) + shape_tail, -1.0)
        own = torch.arange(V) % world == rank
        t[own] = torch.arange(V, dtype=torch.float32)[own].view(-1, *([1] * len(shape_tail))) * (1.0 if not shape_tail else torch.ones(shape_tail))
        return t
    eng = types.SimpleNamespace(table_update_mode="lazy", fm_w_s0=shard(()), fm_w_s1=shard(()) * 2, fm_v_s0=shard((K,)),
                                fm_v_s1=shard((K,)) * 3)
    gather_owner_sharded_slots(eng)
    torch.save({"w0": eng.fm_w_s0, "w1": eng.fm_w_s1, "v0": eng.fm_v_s0, "v1": eng.fm_v_s1}, os.path.join(out_dir, f"g{rank}.pt"))
    dist.destroy_process_group()


def test_owner_sharded_slots_are_gathered_before_a_checkpoint(tmp_path):
    """Multi-GPU lazy update keeps the optimizer slots of row id on rank id % world; before a checkpoint every rank must
    end up with every row (parallel/sync_dp.gather_owner_sharded_slots) -- here over gloo with three ranks."""
    import socket
    import torch
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_gather_proc, args=(3, port, str(tmp_path)), nprocs=3, join=True)
    ref = torch.arange(23, dtype=torch.float32)
    for r in range(3):
        g = torch.load(tmp_path / f"g{r}.pt")
        assert torch.equal(g["w0"], ref) and torch.equal(g["w1"], 2 * ref)
        assert torch.equal(g["v0"], ref[:, None].expand(23, 4)) and torch.equal(g["v1"], 3 * ref[:, None].expand(23, 4))


def _decide_worker(rank, world, port, out_dir):
    from deepfm_b200.parallel.sync_dp import chief_decides
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # the ranks disagree locally in every combination; rank 0's view must win everywhere
    mine = [(False, True)[(i >> rank) & 1] for i in range(4)]
    got = [chief_decides(f, torch.device("cpu")) for f in mine]
    torch.save({"mine": mine, "got": got}, os.path.join(out_dir, f"d{rank}.pt"))
    dist.destroy_process_group()


def test_stop_and_checkpoint_decisions_are_rank_zeros_on_every_rank(tmp_path):
    """Ranks that launch different numbers of fused steps hang in the collective kernels: any clock-based decision is
    rank 0's, broadcast (bench.measure_pipeline, train.save_is_due)."""
    from deepfm_b200.parallel.sync_dp import chief_decides
    assert chief_decides(True, torch.device("cpu")) is True and chief_decides(False, torch.device("cpu")) is False
    mp.spawn(_decide_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    d0, d1 = (torch.load(tmp_path / f"d{i}.pt") for i in range(2))
    assert d0["mine"] != d1["mine"]
    assert d0["got"] == d0["mine"] and d1["got"] == d0["mine"]


def _even_worker(rank, world, port, out_dir, counts, batch):
    from deepfm_b200.parallel.even_stream import EvenStream, control_group
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    path = os.path.join(out_dir, f"ch{rank}.tfrecords")
    with T.TFRecordWriter(path) as w:
        for i in range(counts[rank]):
            w.write_example(float(i & 1), np.arange(8) + (i % 100), np.ones(8, np.float32))
    group = control_group()
    cfg = DeepFMConfig(8, 200, 8, [16, 8], [1.0, 1.0], l2_reg=1e-3)
    tr = Trainer(cfg, TrainerOptions(batch_size=batch, optimizer="GD", learning_rate=0.05, backend="eager", seed=5),
                 device="cpu")
    seen = []
    for epoch in range(2):                       # the wrapper is per epoch, the group is made once
        loader = EvenStream(BatchLoader([path], 8, batch, streaming=True, num_slots=6, num_workers=2,
                                        ids_dtype=torch.int32), group, depth=2)
        n = 0
        for hb in tr.train_batches(loader):      # every step all-reduces gradients over the ranks: uneven counts would hang
            assert hb.count == batch
            n += 1
        loader.close()
        seen.append(n)
    torch.save({"steps": seen, "vars": {k: v.detach().clone() for k, v in tr.variables().items()}},
               os.path.join(out_dir, f"e{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.timeout(240)
@pytest.mark.parametrize("counts", [(10 * 16, 7 * 16 + 5), (3 * 16, 9 * 16), (16, 0), (6 * 16, 4 * 16 + 15, 5 * 16)])
def test_uneven_streams_end_on_the_same_step_everywhere(tmp_path, counts):
    """Pipe mode on several ranks: channels of different length.  Every rank must run exactly as many steps as the
    shortest channel has full batches -- agreed two steps ahead over gloo -- and the replicas must stay identical."""
    batch = 16
    world = len(counts)
    mp.spawn(_even_worker, args=(world, _free_port(), str(tmp_path), counts, batch), nprocs=world, join=True)
    res = [torch.load(tmp_path / f"e{i}.pt") for i in range(world)]
    want = min(c // batch for c in counts)
    for r in res:
        assert r["steps"] == [want, want]
        for k in res[0]["vars"]:
            assert torch.equal(res[0]["vars"][k], r["vars"][k]), k
