"""Asynchronous parameter-server mode on CPU/gloo (reference 1-ps-cpu workflow, SURVEY 3.2 / test plan §4)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.parallel import ps as P
from deepfm_b200.runtime.trainer import Trainer, TrainerOptions
from deepfm_b200.data.loader import HostBatch

CFG = dict(field_size=6, feature_size=120, embedding_size=4, deep_layers=[8, 4], dropout_keep=[1.0, 1.0], l2_reg=1e-3)


def _port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _data(n, B, seed):
    rng = np.random.default_rng(seed)
    return [(torch.from_numpy(rng.integers(1, 120, (B, 6))), torch.from_numpy(rng.random((B, 6)).astype(np.float32)),
             torch.from_numpy((rng.random(B) < 0.3).astype(np.float32))) for _ in range(n)]


def _proc(rank, world, n_ps, port, out_dir, opt):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = DeepFMConfig(**CFG)
    n_workers = world - n_ps
    if rank < n_ps:
        P.ParameterServer(cfg, rank, n_ps, n_workers, opt, 0.05).serve()
    else:
        widx = rank - n_ps
        client = P.PSClient(cfg, n_ps)
        worker = P.PSWorker(cfg, client, seed=7)
        if widx == 0:
            for n, t in worker.model.tf_variables().items():
                client.set(n, t)
            client.broadcast_simple(P.OP_INIT_DONE, aux=0)
        else:
            import time
            while not client.ask(P.OP_IS_INIT):
                time.sleep(0.01)
        for ids, vals, labels in _data(5, 16, seed=100 + widx):
            worker.train_batch(ids, vals, labels)
        client.broadcast_simple(P.OP_DONE)
    dist.destroy_process_group()


def _single_worker_proc(rank, world, port, out_dir, opt):
    """1 PS + 1 worker; the worker saves what it pulls back at the end."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = DeepFMConfig(**CFG)
    if rank == 0:
        P.ParameterServer(cfg, 0, 1, 1, opt, 0.05).serve()
    else:
        client = P.PSClient(cfg, 1)
        worker = P.PSWorker(cfg, client, seed=7)
        for n, t in worker.model.tf_variables().items():
            client.set(n, t)
        client.broadcast_simple(P.OP_INIT_DONE, aux=0)
        for ids, vals, labels in _data(5, 16, seed=100):
            worker.train_batch(ids, vals, labels)
        worker.pull_all()
        torch.save({k: v.detach().clone() for k, v in worker.model.tf_variables().items()}, os.path.join(out_dir, "ps.pt"))
        assert worker.global_step == 5
        client.broadcast_simple(P.OP_DONE)
    dist.destroy_process_group()


@pytest.mark.parametrize("opt", ["Adam", "Adagrad"])
def test_one_worker_ps_equals_local_training(tmp_path, opt):
    """With a single worker the asynchronous PS run is step for step the local run (no LR scaling, PS.py:180)."""
    mp.spawn(_single_worker_proc, args=(2, _port(), str(tmp_path), opt), nprocs=2, join=True)
    got = torch.load(tmp_path / "ps.pt")
    cfg = DeepFMConfig(**CFG)
    tr = Trainer(cfg, TrainerOptions(batch_size=16, optimizer=opt, learning_rate=0.05, backend="eager", seed=7,
                                     scale_lr_by_world=False), device="cpu")
    for ids, vals, labels in _data(5, 16, seed=100):
        tr.train_batch(HostBatch(ids.int(), vals, labels, 16))
    for k, v in tr.variables().items():
        torch.testing.assert_close(got[k], v.detach(), rtol=1e-5, atol=1e-6)


def test_two_ps_four_workers_run_to_completion(tmp_path):
    """BASELINE.json config 4 in miniature: 2 PS + 4 workers, nobody waits for anybody, everybody exits."""
    mp.spawn(_proc, args=(6, 2, _port(), str(tmp_path), "Adam"), nprocs=6, join=True)


def test_variable_placement_is_round_robin_per_variable():
    names = list(DeepFM(DeepFMConfig(**CFG), seed=0).tf_variables().keys())
    place = P.variable_placement(names, 2)
    assert place["fm_bias"] == 0 and place["fm_w"] == 1 and place["fm_v"] == 0      # fm_v whole on one task
    assert sorted(set(place.values())) == [0, 1]


def test_worker_dropout_stream_is_counter_based_and_per_worker():
    """PS workers draw dropout masks from the engine's Philox stream keyed by (dropout_seed, local step, layer): the same
    worker replays the same masks, two workers with different keys do not share them."""
    import torch
    from deepfm_b200.models.deepfm import DeepFMConfig
    from deepfm_b200.parallel import ps as P

    class _NoNet(P.PSClient):                      # a client that answers from a local model: no process group needed
        def __init__(self, cfg):
            super().__init__(cfg, 1)
            from deepfm_b200.models.deepfm import DeepFM
            self.m = DeepFM(cfg, seed=0)
        def pull_step(self, uniq):
            tv = self.m.tf_variables()
            dense = {n: t.detach().clone() for n, t in tv.items() if n not in P.TABLES}
            return dense, tv["fm_w"].detach()[uniq].clone(), tv["fm_v"].detach()[uniq].clone()
        def push_step(self, uniq, gw, gv, values):
            self.dense = {k: v.detach().clone() for k, v in values.items()}
            return 1

    cfg = DeepFMConfig(field_size=6, feature_size=120, embedding_size=4, deep_layers=[16, 8], dropout_keep=[0.5, 0.5])
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(1, 120, (32, 6), generator=g)
    vals, labels = torch.rand(32, 6, generator=g), (torch.rand(32, generator=g) < 0.3).float()

    def grads(dropout_seed):
        c = _NoNet(cfg)
        w = P.PSWorker(cfg, c, seed=0, dropout_seed=dropout_seed)
        w.train_batch(ids, vals, labels)
        return c.dense["Deep-part/mlp0/weights"]

    a, b, c = grads(11), grads(11), grads(12)
    assert torch.equal(a, b) and not torch.equal(a, c)
