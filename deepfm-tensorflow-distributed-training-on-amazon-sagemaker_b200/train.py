"""Task driver: ``python -m deepfm_b200.train --task_type {train,eval,infer,export} ...``

One entry point for what the reference splits over two scripts' ``main`` (HVD.py:331-493,
PS.py:389-551):

* flags of both scripts, environment optional (utils/flags.py);
* file discovery by glob ``**/tr*.tfrecords``, ``**/va*.tfrecords``, ``**/te*.tfrecords``
  (HVD.py:352-362), or FIFO channels in pipe mode (HVD.py:127-149, 420-456);
* rank / shard policy (parallel/env.py, data/sharding.py);
* train: per-epoch train -> evaluate loop (file mode) or one multi-epoch train -> evaluate
  (pipe mode), checkpoints every ``save_checkpoints_secs`` and at the end, resume by default
  from the newest checkpoint, and -- like the reference -- always an export afterwards
  (HVD.py:475);
* eval / infer (``pred.txt`` with one ``%f`` per example, HVD.py:471-473) / export.

Deliberate differences: evaluation covers every record unless
``--drop_remainder_eval``; only rank 0 evaluates, infers and writes while the others wait at a
barrier (Q9, Q10, Q13); every rank shuffles the file list with the same seed (Q12); per-epoch batch
counts are equalised across ranks so no rank waits on a collective that never comes (Q11).
"""
from __future__ import annotations

import glob
import os
import random
import shutil
import sys
import time
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from .data.loader import BatchLoader
from .data.sharding import hvd_shard, ps_shard, shard_batches_per_epoch
from .data.tfrecord import count_records
from .models.deepfm import DeepFMConfig
from .parallel.env import DistEnv, detect_env
from .parallel.even_stream import EvenStream, control_group
from .runtime.trainer import Trainer, TrainerOptions
from .utils.checkpoint import CheckpointManager
from .utils.export import export_servable
from .utils.flags import Flags, channel_names, parse_flags, print_flags


def log(env: DistEnv, *a) -> None:
    print(f"[rank {env.rank}]", *a, flush=True)


def discover_files(flags: Flags):
    if flags.pipe_mode:
        return [], [], []
    tr = sorted(glob.glob(f"{flags.training_data_dir}/**/tr*.tfrecords", recursive=True)) if flags.training_data_dir else []
    va = sorted(glob.glob(f"{flags.val_data_dir}/**/va*.tfrecords", recursive=True)) if flags.val_data_dir else []
    te = sorted(glob.glob(f"{flags.val_data_dir}/**/te*.tfrecords", recursive=True)) if flags.val_data_dir else []
    random.Random(flags.seed).shuffle(tr)       # same order on every rank
    return tr, va, te


def pipe_channel_path(channel: str, epoch: int = 0) -> str:
    """SageMaker pipe mode exposes channel <c> as FIFOs /opt/ml/input/data/<c>_<epoch>."""
    root = os.environ.get("SM_INPUT_DATA_DIR", "/opt/ml/input/data")
    return os.path.join(root, f"{channel}_{epoch}")


def init_distributed(env: DistEnv, flags: Flags) -> None:
    if env.world_size <= 1 or (dist.is_available() and dist.is_initialized()):
        return
    os.environ.setdefault("MASTER_ADDR", env.master_addr)
    os.environ.setdefault("MASTER_PORT", str(env.master_port))
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(env.local_rank % torch.cuda.device_count())
        dist.init_process_group("nccl", rank=env.rank, world_size=env.world_size,
                                device_id=torch.device("cuda", torch.cuda.current_device()))
    else:
        dist.init_process_group("gloo", rank=env.rank, world_size=env.world_size)


def barrier() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def make_loader(flags: Flags, files: Sequence[str], batch_size: int, shard, drop_remainder: bool,
                num_epochs: int = 1, streaming: bool = False, min_slots: int = 2) -> BatchLoader:
    return BatchLoader(list(files), flags.field_size, batch_size, num_epochs=num_epochs, drop_remainder=drop_remainder,
                       shard_n=shard[0], shard_i=shard[1], num_workers=flags.parser_threads,
                       num_slots=max(min_slots, flags.prefetch_batches), ids_dtype=torch.int32, verify_crc=True,
                       streaming=streaming, feature_size=flags.feature_size)


def equalised_batches(env: DistEnv, flags: Flags, files: Sequence[str], shard) -> Optional[int]:
    """Batches per epoch every rank will run (the minimum over ranks), or None when unknown (streams)."""
    if not files:
        return None
    n = sum(count_records(f, verify_crc=False) for f in files)
    mine = shard_batches_per_epoch(n, shard[0], shard[1], flags.batch_size, drop_remainder=True)
    if env.world_size > 1:
        t = torch.tensor([mine], dtype=torch.int64, device="cuda" if torch.cuda.is_available() else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        mine = int(t.item())
    return mine


def run_eval(trainer: Trainer, flags: Flags, files: Sequence[str], streaming: bool = False) -> dict:
    loader = make_loader(flags, files, flags.batch_size, (1, 0), drop_remainder=flags.drop_remainder_eval,
                         streaming=streaming)
    try:
        return trainer.evaluate(loader, auc_mode=flags.auc_mode)
    finally:
        loader.close()


def save(trainer: Trainer, ckpt: CheckpointManager, env: DistEnv) -> None:
    """COLLECTIVE: every rank drains its queue and waits on the host while the chief writes, so no rank sits inside a
    collective kernel waiting for a chief that is busy with the file system."""
    trainer.prepare_checkpoint()
    if env.is_chief and ckpt.model_dir:
        p = ckpt.save(trainer.global_step, trainer.checkpoint_tensors())
        log(env, f"Saving checkpoints for {trainer.global_step} into {p}")
    barrier()


def save_is_due(trainer: Trainer, ckpt: CheckpointManager, env: DistEnv, every: int = 64) -> bool:
    """The chief's timer decides, for everybody: asked every `every` steps with one tiny broadcast."""
    if env.world_size <= 1:
        return ckpt.due()
    if trainer.global_step % every:
        return False
    t = torch.tensor([1 if (env.is_chief and ckpt.due()) else 0], dtype=torch.int32,
                     device="cuda" if torch.cuda.is_available() else "cpu")
    dist.broadcast(t, src=0)
    return bool(t.item())


def train_epochs(trainer: Trainer, flags: Flags, env: DistEnv, files: Sequence[str], shard, ckpt: CheckpointManager,
                 num_epochs: int, streaming: bool, max_batches: Optional[int], even_group=None) -> None:
    """``even_group`` (a gloo group over all ranks): the batch count is not known up front (streams), so the ranks
    agree step by step, two steps ahead, on whether everybody still has a batch (parallel/even_stream.py)."""
    loader = make_loader(flags, files, flags.batch_size, shard, drop_remainder=True, num_epochs=num_epochs,
                         streaming=streaming, min_slots=6 if even_group is not None else 2)
    if even_group is not None:
        loader = EvenStream(loader, even_group, depth=2)
    t0, s0 = time.time(), trainer.global_step
    limit = None if max_batches is None else max_batches * num_epochs
    done = 0
    try:
        for hb in trainer.train_batches(loader, limit=limit):
            done += 1
            if flags.log_steps > 0 and trainer.global_step % flags.log_steps == 0 and env.is_chief:
                dt = max(time.time() - t0, 1e-9)
                sps = (trainer.global_step - s0) / dt
                log(env, f"loss = {trainer.last_loss():.6f}, step = {trainer.global_step} "
                         f"({sps:.1f} global_step/sec, {sps * flags.batch_size * env.world_size:.0f} examples/sec)")
                t0, s0 = time.time(), trainer.global_step
            if save_is_due(trainer, ckpt, env):
                save(trainer, ckpt, env)
            if flags.max_steps and trainer.global_step >= flags.max_steps:
                break
    finally:
        loader.close()


def build_trainer(flags: Flags, env: DistEnv, scale_lr: bool) -> Trainer:
    cfg = DeepFMConfig(field_size=flags.field_size, feature_size=flags.feature_size,
                       embedding_size=flags.embedding_size, deep_layers=flags.deep_layer_list,
                       dropout_keep=flags.dropout_keep_list, batch_norm=flags.batch_norm,
                       batch_norm_decay=flags.batch_norm_decay, l2_reg=flags.l2_reg, loss_type=flags.loss_type)
    opts = TrainerOptions(batch_size=flags.batch_size, optimizer=flags.optimizer, learning_rate=flags.learning_rate,
                          table_update_mode=flags.table_update_mode, backend=flags.backend, seed=flags.seed,
                          cuda_graph=flags.cuda_graph, scale_lr_by_world=scale_lr, log_steps=flags.log_steps)
    device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    return Trainer(cfg, opts, device=device)


def main(argv: Optional[List[str]] = None) -> int:
    flags, ignored = parse_flags(argv)
    env = detect_env()
    mode = flags.dist
    if mode == "auto":
        mode = "ps" if (flags.job_name or os.environ.get("TF_CONFIG")) and flags.dist_mode else \
               ("hvd" if env.world_size > 1 else "none")
    if mode == "ps":
        from .parallel.ps import run_ps
        return run_ps(flags, env)

    init_distributed(env, flags)
    if env.is_chief:
        print_flags(flags)
        if ignored:
            print("ignored unknown arguments:", " ".join(ignored))
    tr_files, va_files, te_files = discover_files(flags)
    log(env, "tr_files:", tr_files)
    log(env, "va_files:", va_files)
    log(env, "te_files:", te_files)

    model_dir = flags.effective_model_dir
    if flags.clear_existing_model and model_dir and env.is_chief:      # rank 0 only
        try:
            shutil.rmtree(model_dir)
            log(env, f"existing model cleaned at {model_dir}")
        except OSError as e:
            log(env, e, "at clear_existing_model")
    barrier()

    log(env, f"current rank is {env.rank} of {env.world_size}; hosts {flags.hosts}; current host {flags.current_host}")
    trainer = build_trainer(flags, env, scale_lr=True)
    ckpt = CheckpointManager(model_dir if env.is_chief else "", keep_max=flags.keep_checkpoint_max,
                             save_secs=flags.save_checkpoints_secs)
    # resume by default from the newest checkpoint (Estimator behaviour), then broadcast from rank 0
    if model_dir:
        restored = CheckpointManager(model_dir, flags.keep_checkpoint_max).restore_latest() if env.is_chief else None
        if restored is not None:
            step, tensors = restored
            trainer.load_checkpoint_tensors(step, tensors)
            log(env, f"Restoring parameters from {model_dir} at step {step}")
        trainer.broadcast_parameters()

    channels = channel_names()
    eval_channel = flags.evaluation_channel_name or (channels[0] if channels else "evaluation")
    shard = hvd_shard(flags.pipe_mode, flags.enable_s3_shard, flags.enable_data_multi_path, flags.worker_per_host,
                      len(flags.hosts), env.rank, env.local_rank, env.world_size)

    if flags.task_type == "train":
        if not flags.pipe_mode:
            if not tr_files:
                raise SystemExit(f"no tr*.tfrecords under --training_data_dir {flags.training_data_dir!r}")
            nb = equalised_batches(env, flags, tr_files, shard)
            for epoch in range(flags.num_epochs):
                train_epochs(trainer, flags, env, tr_files, shard, ckpt, 1, False, nb)
                save(trainer, ckpt, env)
                if env.is_chief and va_files:
                    log(env, f"epoch {epoch + 1}: " + str(run_eval(trainer, flags, va_files)))
                barrier()
                if flags.max_steps and trainer.global_step >= flags.max_steps:
                    break
        else:
            # one channel per local worker after the evaluation channel (HVD.py:445, README.md:79-83)
            ch = flags.training_channel_name or (channels[1 + env.local_rank] if len(channels) > 1 + env.local_rank
                                                 else f"training_{env.local_rank}")
            paths = [pipe_channel_path(ch, e) for e in range(flags.num_epochs)]
            # channels cannot be counted and rarely hold the same number of records: every epoch ends, on every rank,
            # at the first step some rank has no full batch for (the reference hangs in Horovod there)
            group = control_group() if env.world_size > 1 else None
            for e, pth in enumerate(paths):
                train_epochs(trainer, flags, env, [pth], shard, ckpt, 1, True, None, even_group=group)
            save(trainer, ckpt, env)
            if env.is_chief:
                log(env, str(run_eval(trainer, flags, [pipe_channel_path(eval_channel, 0)], streaming=True)))
            barrier()
    elif flags.task_type == "eval":
        if env.is_chief:
            files = va_files if not flags.pipe_mode else [pipe_channel_path(eval_channel, 0)]
            log(env, str(run_eval(trainer, flags, files, streaming=bool(flags.pipe_mode))))
        barrier()
    elif flags.task_type == "infer":
        if env.is_chief:
            loader = make_loader(flags, te_files, flags.batch_size, (1, 0), drop_remainder=flags.drop_remainder_eval)
            out = os.path.join(flags.val_data_dir, "pred.txt")
            with open(out, "w") as fo:
                for probs in trainer.predict(loader):
                    for p in probs:
                        fo.write("%f\n" % p)
            loader.close()
            log(env, f"predictions written to {out}")
        barrier()

    if flags.task_type in ("export", "train") and flags.servable_model_dir:
        if env.is_chief:
            path = export_servable(trainer.export_model(), flags.servable_model_dir)
            log(env, f"servable model exported to {path}")
        barrier()
    trainer.close()
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
