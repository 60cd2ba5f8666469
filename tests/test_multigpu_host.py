"""CPU test (gloo, world_size 2) of the N>1 host logic: contiguous sharding, rank bases, and the contract of the
merge — all-gathered per-rank sorted runs merged by (bucket, time key, rank) equal the whole-job dequeue order.
The per-rank engine is played by the oracle here; the GPU merge kernels (concat_kernel, merge_strided_kernel) are checked
against the same reference merge and against the whole-job oracle order in tests/test_gpu_multigpu.py (two GPUs)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import llm_message_queue_b200  # noqa: F401  (registers the package so the submodule import below works)
from llm_message_queue_b200 import multigpu
from oracle import c_oracle as co

N_TOTAL = 6001


def reference_merge(runs):
    """runs: list over ranks of (bucket u8[], key u64[], gidx u32[]) each sorted by (bucket, key); stable by rank."""
    b = np.concatenate([r[0] for r in runs]).astype(np.int64)
    k = np.concatenate([r[1] for r in runs])
    g = np.concatenate([r[2] for r in runs])
    rk = np.concatenate([np.full(len(r[0]), i) for i, r in enumerate(runs)])
    pos = np.concatenate([np.arange(len(r[0])) for r in runs])
    order = np.lexsort((pos, rk, k, b))
    return g[order]


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        off, data, meta, ts = co.generate(0xBEEF, N_TOTAL, 32, 700, edge=100)
        bounds = multigpu.shard_bounds(N_TOTAL, world)
        lo, hi = int(bounds[rank]), int(bounds[rank + 1])
        so = (off[lo:hi + 1] - off[lo]).astype(np.uint64)
        sd = data[int(off[lo]):int(off[hi])]
        res, order = co.pipeline(co.Dict(), so, sd, meta[lo:hi], ts[lo:hi], rules_on=True)
        bucket = multigpu.bucket_of(res)
        run = (bucket[order], (multigpu.rank_bases(bounds)[rank] + order).astype(np.uint64),
               (int(bounds[rank]) + order).astype(np.uint32))
        gathered = [None] * world
        dist.all_gather_object(gathered, run)
        merged = reference_merge(gathered)
        if rank == 0:
            whole_res, whole_order = co.pipeline(co.Dict(), off, data, meta, ts, rules_on=True)
            out.put((np.array_equal(merged, whole_order), len(merged), len(whole_order)))
    finally:
        dist.destroy_process_group()


def test_sharded_runs_merge_to_the_whole_job_order_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, a, b = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok and a == b


def test_shard_bounds():
    b = multigpu.shard_bounds(10, 4)
    assert b.tolist() == [0, 3, 6, 8, 10]
    assert multigpu.rank_bases(b).tolist() == [0, 3, 6, 8]
    assert multigpu.shard_bounds(0, 2).tolist() == [0, 0, 0]
