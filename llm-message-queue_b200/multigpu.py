"""Host-side plumbing of the multi-GPU path: one process per GPU, messages sharded by contiguous ranges of
the submit sequence, torch.distributed only carries the NCCL unique id and the barrier.  The exchange itself
(all-gather of sorted runs + merge) is triage_merge_allgather in the CUDA library."""
from __future__ import annotations

import numpy as np


def shard_bounds(n_total: int, world: int) -> np.ndarray:
    """rank r owns global messages [b[r], b[r+1]) — contiguous, sizes differ by at most one."""
    base, rem = divmod(n_total, world)
    sizes = np.array([base + (1 if r < rem else 0) for r in range(world)], dtype=np.uint64)
    b = np.zeros(world + 1, dtype=np.uint64)
    np.cumsum(sizes, out=b[1:])
    return b


def rank_bases(bounds: np.ndarray) -> np.ndarray:
    """what triage_merge_allgather takes as rank_base: first global index of every rank"""
    return np.ascontiguousarray(bounds[:-1], dtype=np.uint64)


def bucket_of(results: np.ndarray) -> np.ndarray:
    """bucket byte the device orders by: (queue level - 1) * 4 + (item priority - 1), 255 = not admitted"""
    b = (results["queue_level"].astype(np.int64) - 1) * 4 + (results["label"].astype(np.int64) - 1)
    return np.where(results["accepted"] != 0, b, 255).astype(np.uint8)


def init_comm(engine, dist) -> None:
    """rank 0 creates the NCCL id, everybody joins (dist = an initialised torch.distributed module)."""
    from ._capi import Engine
    rank, world = dist.get_rank(), dist.get_world_size()
    uid = [Engine.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    engine.comm_init(world, rank, uid[0])
