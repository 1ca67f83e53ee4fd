"""Record-level shard policies of the reference, as one pure function.

All policies are ``dataset.shard(n, i)`` applied per record before batching (HVD.py:130-151,
PS.py:153-156; matrix in README.md:87-92):

    HVD file mode, S3-sharded data        (worker_per_host, local_rank)      HVD.py:130-131
    HVD file mode, fully replicated data  (world_size, rank)                 HVD.py:132-133
    HVD pipe, multi path + S3 shard       no shard                           HVD.py:140-141
    HVD pipe, multi path, no S3 shard     (num_hosts, rank // worker_per_host) if hosts > 1   HVD.py:142-144
    HVD pipe, single path + S3 shard      (worker_per_host, local_rank)      HVD.py:146-147
    HVD pipe, single path, no S3 shard    (world_size, rank)                 HVD.py:148-149
    PS, not S3-sharded                    (num_hosts, index of current host) PS.py:153-156
    PS, S3-sharded                        no shard                           PS.py:153
"""
from __future__ import annotations

from typing import Sequence, Tuple


def hvd_shard(pipe_mode: int, enable_s3_shard: bool, enable_data_multi_path: bool, worker_per_host: int,
              num_hosts: int, rank: int, local_rank: int, world_size: int) -> Tuple[int, int]:
    """Returns (n, i): keep records whose stream index % n == i.  (1, 0) means no sharding."""
    if not pipe_mode:
        if enable_s3_shard:
            return max(1, worker_per_host), local_rank
        return max(1, world_size), rank
    if enable_data_multi_path:
        if not enable_s3_shard and num_hosts > 1:
            return num_hosts, rank // max(1, worker_per_host)
        return 1, 0
    if enable_s3_shard:
        return max(1, worker_per_host), local_rank
    return max(1, world_size), rank


def ps_shard(enable_s3_shard: bool, hosts: Sequence[str], current_host: str) -> Tuple[int, int]:
    if enable_s3_shard or len(hosts) <= 1:
        return 1, 0
    return len(hosts), list(hosts).index(current_host)


def shard_batches_per_epoch(n_records: int, shard_n: int, shard_i: int, batch_size: int,
                            drop_remainder: bool = True) -> int:
    """Batches one rank gets from a stream of ``n_records`` (ranks can differ by one: SURVEY Q11)."""
    mine = (n_records - shard_i + shard_n - 1) // shard_n if n_records > shard_i else 0
    return mine // batch_size if drop_remainder else (mine + batch_size - 1) // batch_size
