"""Equal step counts over streams of unknown length (SageMaker pipe mode on several ranks).

In file mode every rank can count its records, so the per-epoch batch count is equalised up front with one
all-reduce (``train.equalised_batches``).  A FIFO cannot be counted: each rank finds out that its channel
is exhausted only when it gets there, and the channels of a job rarely hold the same number of records.  With
synchronous data parallelism a rank that runs one step more than a peer waits for that peer forever -- Horovod hangs
in its all-reduce there (HVD.py:133,151,296), and the fused collective kernels here would sit in their flag waits until
the watchdog traps.

:class:`EvenStream` wraps a batch loader and ends the stream on EVERY rank at the first step for which SOME rank has
no full batch.  The agreement costs nothing on the step path: whether a rank has batch ``s`` is known as soon as
its loader delivered it, which is ``depth`` steps before it is trained on, and is reduced over the ranks (MIN) with an
asynchronous all-reduce on a *gloo* group -- host memory, no CUDA stream involved -- whose result is only read when
step ``s`` is due.  Every rank posts the reductions for steps ``0, 1, 2, ...`` in order, and has posted exactly those
up to ``s + depth`` when it reads the one for ``s``, so all ranks stop at the same ``s`` with the same number of
collectives issued.
"""
from __future__ import annotations

from collections import deque
from typing import Deque, Optional, Tuple

import torch
import torch.distributed as dist


def control_group():
    """COLLECTIVE: a gloo group over all ranks for host-side agreement (the default group may be NCCL)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() < 2:
        return None
    try:
        return dist.new_group(backend="gloo")
    except Exception as e:       # noqa: BLE001 -- e.g. a host name gloo cannot resolve: the same on every rank of a box
        if dist.get_rank() == 0:
            print(f"[even_stream] no gloo control group ({type(e).__name__}: {e}); channels of unequal length will "
                  "not be evened out", flush=True)
        return None


class EvenStream:
    """Iterates ``loader`` while every rank of ``group`` still has a FULL batch; see the module docstring.

    Looks like a :class:`~deepfm_b200.data.loader.BatchLoader` to ``Trainer.train_batches`` (``defer_release``,
    ``num_slots``, ``release``).  It keeps ``depth`` batches of look-ahead out of the loader's ring, so the ring must have
    at least ``depth + 4`` slots for the pipelined trainer."""

    def __init__(self, loader, group, depth: int = 2, full_batch: Optional[int] = None):
        if getattr(loader, "num_slots", depth + 2) < depth + 2:
            raise ValueError(f"EvenStream(depth={depth}) needs a loader ring of at least {depth + 2} slots")
        self.loader, self.group, self.depth = loader, group, max(1, depth)
        self.full = full_batch if full_batch is not None else loader.batch_size
        self.num_slots = max(2, getattr(loader, "num_slots", 4) - self.depth)     # what is left for the consumer
        self.defer_release = False           # set by Trainer.train_batches: the consumer returns slots itself
        loader.defer_release = True          # this wrapper always holds several batches of the loader
        self._it = iter(loader)
        self._ahead: Deque[Tuple[object, object, torch.Tensor]] = deque()    # (batch | None, work, flag)
        self._eof = False
        self._last_slot: Optional[int] = None
        self._done = False
        self.steps_agreed = 0

    # -- BatchLoader surface ------------------------------------------------------------------------------
    @property
    def batch_size(self):
        return self.loader.batch_size

    def release(self, slot: int) -> None:
        self.loader.release(slot)

    def stats(self):
        return self.loader.stats()

    def close(self) -> None:
        self._drain()
        self.loader.close()

    def __iter__(self):
        return self

    # -- the agreement --------------------------------------------------------------------------------------
    def _post_next(self) -> None:
        hb = None
        if not self._eof:
            hb = next(self._it, None)
            if hb is None or hb.count != self.full:       # a short tail batch does not count: drop_remainder semantics
                if hb is not None and hb.slot is not None:
                    self.loader.release(hb.slot)
                hb, self._eof = None, True
        flag = torch.tensor([1 if hb is not None else 0], dtype=torch.int32)
        work = dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group, async_op=True) if self.group is not None else None
        self._ahead.append((hb, work, flag))

    def __next__(self):
        if self._done:
            raise StopIteration
        if self._last_slot is not None and not self.defer_release:
            self.loader.release(self._last_slot)          # plain iteration: a batch is valid until the next one is asked for
        self._last_slot = None
        while len(self._ahead) <= self.depth:             # know (and announce) steps s .. s + depth
            self._post_next()
        hb, work, flag = self._ahead.popleft()
        if work is not None:
            work.wait()
        if int(flag.item()) == 0:                         # some rank is out of data at this step: everybody stops here
            if hb is not None and hb.slot is not None:
                self.loader.release(hb.slot)
            self._done = True
            self._drain()
            raise StopIteration
        self.steps_agreed += 1
        self._last_slot = hb.slot
        return hb

    def _drain(self) -> None:
        """Completes the reductions already posted (every rank has posted the same ones) and returns their slots."""
        while self._ahead:
            hb, work, _ = self._ahead.popleft()
            if work is not None:
                work.wait()
            if hb is not None and hb.slot is not None:
                self.loader.release(hb.slot)
        if self._last_slot is not None and not self.defer_release:
            self.loader.release(self._last_slot)
            self._last_slot = None
