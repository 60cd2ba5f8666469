"""GPU test of triage_submit_device_static — the submit for callers that must not allocate device memory
(serving loops that cannot take cudaMalloc's implicit synchronisation, hosts that own all of HBM): it has to give
exactly what triage_submit_device gives, on back-to-back calls that share one fixed scratch array, and it has to turn
away batches beyond its stated limits instead of allocating.  Collected last on purpose (file name), and self-contained
so it also passes when run alone in a fresh process."""
import numpy as np
import pytest

import llm_message_queue_b200 as mq
from oracle import c_oracle as co

pytestmark = pytest.mark.gpu


def _upload(torch, dev, off, data, meta, ts):
    n = len(meta)
    pad = np.zeros(((len(data) + 31) // 16) * 16, np.uint8)
    pad[:len(data)] = data
    return dict(n=n, total=len(data),
                bytes=torch.from_numpy(pad).to(dev), off=torch.from_numpy(off.view(np.int64)).to(dev),
                meta=torch.from_numpy(meta.view(np.uint8).reshape(n, 16)).to(dev),
                ts=torch.from_numpy(ts.view(np.int64)).to(dev),
                res=torch.zeros((n, 16), dtype=torch.uint8, device=dev), order=torch.zeros(n, dtype=torch.int32, device=dev))


def _run(eng, b, static, use_ts):
    fn = eng.submit_device_static if static else eng.submit_device
    b["res"].zero_(); b["order"].fill_(-1)
    adm = fn(b["off"].data_ptr(), b["bytes"].data_ptr(), b["total"], b["meta"].data_ptr(), b["ts"].data_ptr() if use_ts else None,
             b["n"], b["res"].data_ptr(), b["order"].data_ptr())
    return adm, b["res"].cpu().numpy().copy(), b["order"].cpu().numpy()[:adm].copy()


def test_static_scratch_submit_matches_the_allocating_one_twice_in_a_row():
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda:0")
    eng = mq.Engine(device=0, rules_enabled=True, max_queue_size=4000)
    try:
        eng.commit()    # the no-allocation entry point refuses to compile the dictionary itself
        big = _upload(torch, dev, *co.generate(71, 30000, 64, 3000, edge=100))
        small = _upload(torch, dev, *co.generate(72, 900, 1, 200, edge=2000))
        rng = np.random.default_rng(3)
        shuffled = _upload(torch, dev, *co.generate(73, 5000, 16, 400, edge=50))
        shuffled["ts"] = torch.from_numpy(rng.permutation(5000).astype(np.int64) * 7 + 10**18).to(dev)
        torch.cuda.synchronize()
        # two calls in a row on the shared array (a big batch, then a smaller one that must not see stale counters),
        # then a batch whose enqueue times are out of order (radix passes use the array's permutation buffers)
        for batch, use_ts in ((big, False), (small, True), (big, True), (shuffled, True)):
            want = _run(eng, batch, static=False, use_ts=use_ts)
            for _ in range(2):
                got = _run(eng, batch, static=True, use_ts=use_ts)
                assert got[0] == want[0]
                assert np.array_equal(got[1], want[1])      # labels, analysis, admission: bit-exact
                assert np.array_equal(got[2], want[2])      # dequeue order: bit-exact
        # beyond the stated maximum: refused with a clear message, nothing allocated on its behalf
        with pytest.raises(mq.TriageError) as ei:
            eng.submit_device_static(big["off"].data_ptr(), big["bytes"].data_ptr(), big["total"], big["meta"].data_ptr(), None,
                                     65537, big["res"].data_ptr(), big["order"].data_ptr())
        assert ei.value.status == mq._capi.E_ARG and "65536" in ei.value.message
        with pytest.raises(mq.TriageError) as ei:
            eng.submit_device_static(big["off"].data_ptr(), big["bytes"].data_ptr(), (256 << 20) + 1, big["meta"].data_ptr(), None,
                                     10, big["res"].data_ptr(), big["order"].data_ptr())
        assert ei.value.status == mq._capi.E_ARG and "exceeds" in ei.value.message
        # an engine whose dictionary was never committed is told to commit first
        fresh = mq.Engine(device=0)
        try:
            with pytest.raises(mq.TriageError) as ei:
                fresh.submit_device_static(small["off"].data_ptr(), small["bytes"].data_ptr(), small["total"], small["meta"].data_ptr(),
                                           None, small["n"], small["res"].data_ptr(), small["order"].data_ptr())
            assert "triage_dict_commit" in ei.value.message
        finally:
            fresh.close()
    finally:
        eng.close()
