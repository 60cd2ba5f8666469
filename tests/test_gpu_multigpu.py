"""GPU test of the multi-GPU exchange (needs two visible GPUs, skipped otherwise): two ranks, one engine per GPU,
triage_submit_device_padded on each shard, then triage_merge_allgather.  Both merge kernels are driven -
contiguous sequence shards take the offset-arithmetic merge (concat_kernel), explicit enqueue times that interleave
across the ranks take the ranked merge (merge_strided_kernel) - and the merged whole-job dequeue order on every rank
is compared with (a) the CPU oracle run on the whole job as one batch and (b) the reference merge of
tests/test_multigpu_host.py over the per-rank runs.  Ranks admit different numbers of messages (explicit priorities
that name no queue are turned away) and most of the 16 (level, priority) buckets are empty."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N_PER_RANK = 20011


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    import llm_message_queue_b200 as mq
    from llm_message_queue_b200 import multigpu
    from oracle import c_oracle as co
    from tests.test_multigpu_host import reference_merge
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        n = N_PER_RANK
        # the whole job, generated identically on every rank; rank r owns the contiguous range [r*n, (r+1)*n)
        off, data, meta, ts = co.generate(0x2D0C, n * world, 16, 900, edge=300)
        # an explicit priority that names no queue is turned away (PushMessage: queue not found): rank r loses 37 * (r + 1)
        for r in range(world):
            meta["preset"][r * n + 11:r * n + 11 + 37 * (r + 1) * 3:3] = 7
        lo, hi = rank * n, (rank + 1) * n
        soff = (off[lo:hi + 1] - off[lo]).astype(np.uint64)
        sdata = data[int(off[lo]):int(off[hi])]
        eng = mq.Engine(device=rank)
        multigpu.init_comm(eng, dist)
        poff, plen, pbytes = eng.pack_padded(soff, sdata)
        d_bytes = torch.from_numpy(pbytes).to(dev)
        d_len = torch.from_numpy(plen.view(np.int32)).to(dev)
        d_off = torch.from_numpy(poff.view(np.int64)).to(dev)
        d_meta = torch.from_numpy(meta[lo:hi].view(np.uint8).reshape(n, 16)).to(dev)
        d_res = torch.zeros((n, 16), dtype=torch.uint8, device=dev)
        d_ord = torch.zeros(n, dtype=torch.int32, device=dev)
        d_gord = torch.zeros(n * world, dtype=torch.int32, device=dev)
        rank_base = np.arange(world, dtype=np.uint64) * np.uint64(n)
        results = {}
        for case in ("contiguous", "contiguous-async", "contiguous-two-step", "interleaved"):
            os.environ.pop("TRIAGE_MERGE_TWO_STEP", None)
            if case == "contiguous-two-step":          # the header + host geometry + order exchange kept for explicit times
                os.environ["TRIAGE_MERGE_TWO_STEP"] = "1"
            if case.startswith("contiguous"):
                ts_all = ts                                   # generator times: strictly increasing in global arrival order
                ts_ptr = None                                 # engine-stamped: arrival order is time order
            else:                                             # rank r's message i gets a time between those of the other ranks
                ts_all = np.empty(n * world, np.uint64)
                for r in range(world):
                    ts_all[r * n:(r + 1) * n] = (np.arange(n, dtype=np.uint64) * np.uint64(world) + np.uint64(r)) * np.uint64(10) + np.uint64(10**18)
                d_ts = torch.from_numpy(ts_all[lo:hi].view(np.int64).copy()).to(dev)
                ts_ptr = d_ts.data_ptr()
            torch.cuda.synchronize()
            adm = eng.submit_device_padded(d_off.data_ptr(), d_len.data_ptr(), d_bytes.data_ptr(), int(poff[-1]), d_meta.data_ptr(), ts_ptr,
                                           n, d_res.data_ptr(), d_ord.data_ptr())
            d_gord.fill_(-1)
            if case == "contiguous-async":             # one exchange on the merge stream, a second submit issued under it
                eng.merge_allgather_async(rank_base, d_gord.data_ptr(), n * world)
                adm2 = eng.submit_device_padded(d_off.data_ptr(), d_len.data_ptr(), d_bytes.data_ptr(), int(poff[-1]), d_meta.data_ptr(), ts_ptr,
                                                n, d_res.data_ptr(), d_ord.data_ptr())
                assert adm2 == adm
                total = eng.merge_wait()
            else:
                total = eng.merge_allgather(rank_base, d_gord.data_ptr(), n * world)
            got = d_gord.cpu().numpy()[:total].astype(np.uint32)
            # (a) the whole job through the oracle as ONE batch
            exp, order = co.pipeline(co.Dict(), off, data, meta, ts_all, nthreads=4)
            # (b) the reference merge of the per-rank runs (each rank's own order from the device)
            res = d_res.cpu().numpy().view(mq.RESULT_DTYPE).reshape(n)
            bucket = multigpu.bucket_of(res)
            loc = d_ord.cpu().numpy()[:adm].astype(np.int64)
            run = (bucket[loc], ts_all[lo:hi][loc] if case == "interleaved" else (rank_base[rank] + loc.astype(np.uint64)),
                   (lo + loc).astype(np.uint32))
            gathered = [None] * world
            dist.all_gather_object(gathered, run)
            merged = reference_merge(gathered)
            adms = [None] * world
            dist.all_gather_object(adms, adm)
            results[case] = (bool(np.array_equal(got, order)), bool(np.array_equal(got, merged)), int(total), int(len(order)), adms)
        eng.close()
        q.put((rank, results))
    finally:
        dist.destroy_process_group()


def test_two_gpu_merge_both_kernels_against_the_oracle():
    torch = pytest.importorskip("torch")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, results in got:
        for case, (vs_oracle, vs_merge, total, want_total, adms) in results.items():
            assert total == want_total, (rank, case, total, want_total)
            assert vs_oracle, (rank, case, "merged order differs from the oracle's whole-job order")
            assert vs_merge, (rank, case, "merged order differs from the reference merge of the per-rank runs")
            assert len(set(adms)) == 2, "the ranks should admit different numbers of messages"
