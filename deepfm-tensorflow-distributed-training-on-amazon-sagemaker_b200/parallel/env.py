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
