"""Who am I?  Rank discovery without MPI.

The reference asks Horovod (``hvd.rank()/local_rank()/size()``, HVD.py:131-149,395,402) after
``hvd.init()`` (HVD.py:333) brought up MPI.  Here the launcher is ``torchrun`` (or mpirun, or
SageMaker), and the same three numbers come from the environment, in this order:

    torchrun        RANK / LOCAL_RANK / WORLD_SIZE / LOCAL_WORLD_SIZE
    Open MPI        OMPI_COMM_WORLD_RANK / _LOCAL_RANK / _SIZE / _LOCAL_SIZE
    nothing set     single process: rank 0 of 1
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional


@dataclass(frozen=True)
class DistEnv:
    rank: int = 0
    local_rank: int = 0
    world_size: int = 1
    local_world_size: int = 1
    master_addr: str = "127.0.0.1"
    master_port: int = 29500

    @property
    def is_chief(self) -> bool:
        return self.rank == 0

    @property
    def distributed(self) -> bool:
        return self.world_size > 1


def _first_int(names: List[str], default: int) -> int:
    for n in names:
        v = os.environ.get(n)
        if v not in (None, ""):
            try:
                return int(v)
            except ValueError:
                pass
    return default


def detect_env() -> DistEnv:
    world = _first_int(["WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE"], 1)
    rank = _first_int(["RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK"], 0)
    local_rank = _first_int(["LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK"], rank)
    local_world = _first_int(["LOCAL_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_SIZE"], world)
    return DistEnv(rank=rank, local_rank=local_rank, world_size=world, local_world_size=local_world,
                   master_addr=os.environ.get("MASTER_ADDR", "127.0.0.1"),
                   master_port=_first_int(["MASTER_PORT"], 29500))


def set_dist_env(dist_mode: int, ps_hosts: str, worker_hosts: str, job_name: str, task_index: int,
                 chief_hosts: str = "") -> Optional[dict]:
    """Builds (and exports) ``TF_CONFIG`` from the --ps_hosts/--worker_hosts/--job_name/--task_index flags.

    The reference ships this helper as dead code with two bugs (PS.py:341-386, SURVEY C9: mode 1 reads an
    undefined ``FLAGS.chief_hosts``, mode 2 drops a worker); this is what it was evidently meant to do:

    * ``dist_mode == 1``: cluster = {chief, ps} (+ workers when given); a ``worker`` task keeps its index.
    * ``dist_mode == 2``: worker 0 becomes the chief, worker 1 the evaluator, workers 2.. become workers 0..;
      ``ps`` tasks keep their index.

    Returns the dict that was written to ``os.environ['TF_CONFIG']`` (None when ``dist_mode`` is 0)."""
    import json
    ps = [h for h in str(ps_hosts).split(",") if h]
    workers = [h for h in str(worker_hosts).split(",") if h]
    if dist_mode == 1:
        chief = [h for h in str(chief_hosts).split(",") if h] or workers[:1]
        rest = workers[1:] if not chief_hosts else workers
        cluster = {"chief": chief, "ps": ps}
        if rest:
            cluster["worker"] = rest
        if job_name == "worker" and not chief_hosts:
            task = {"type": "chief", "index": 0} if task_index == 0 else {"type": "worker", "index": task_index - 1}
        else:
            task = {"type": job_name, "index": task_index}
    elif dist_mode == 2:
        cluster = {"chief": workers[:1], "ps": ps}
        if len(workers) > 2:
            cluster["worker"] = workers[2:]
        if job_name == "worker":
            task = ({"type": "chief", "index": 0} if task_index == 0 else
                    {"type": "evaluator", "index": 0} if task_index == 1 else
                    {"type": "worker", "index": task_index - 2})
        else:
            task = {"type": job_name, "index": task_index}
    else:
        return None
    tf_config = {"cluster": cluster, "task": task}
    os.environ["TF_CONFIG"] = json.dumps(tf_config)
    return tf_config
