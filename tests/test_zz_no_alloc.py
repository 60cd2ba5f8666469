"""GPU test of triage_submit_device_static — the submit for callers that must not allocate device memory
(serving loops that cannot take cudaMalloc's implicit synchronisation, hosts that own all of HBM): it has to give
exactly what triage_submit_device gives, on back-to-back calls that share one fixed scratch array, and it has to turn
away batches beyond its stated limits instead of allocating.  Collected last on purpose (file name), and self-contained
so it also passes when run alone in a fresh process."""
import os

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


def _oracle_inputs(seed, n, lo, hi, edge, extra=()):
    off, data, meta, ts = co.generate(seed, n, lo, hi, edge=edge)
    if extra:   # sprinkle the self-overlapping literals so the sequential recount (fixup_kernel) has work
        rng = np.random.default_rng(seed)
        contents = [bytes(data[int(off[i]):int(off[i + 1])]) for i in range(n)]
        for k in range(0, n, 5):
            unit = extra[k % len(extra)]
            pos = int(rng.integers(0, len(contents[k]) + 1))
            contents[k] = contents[k][:pos] + unit * int(rng.integers(1, 40)) + contents[k][pos:]
        off, data = co.pack(contents)
    return off, data, meta, ts


def test_static_scratch_submit_against_the_oracle_allocates_nothing():
    """The entry point's own promise, checked directly: on an engine that has never run the allocating submit, results
    equal the CPU oracle's (not just the other entry point's), a dictionary with self-overlapping literals goes through
    the recount kernel on the static scratch too, and the free device memory does not move across the calls - a caller
    that owns all of HBM would see a cudaMalloc here as an out-of-memory error."""
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda:0")
    extra = [(4, False, b"haha"), (2, True, b"aa"), (1, False, b"xyxyx")]
    d = co.Dict()
    eng = mq.Engine(device=0, rules_enabled=True, max_queue_size=2500)
    try:
        for p, ci, lit in extra:
            assert d.add(p, lit, ci) == 0
            eng.add_literal(p, lit, ci)
        eng.commit()
        cases = []
        for seed, n, lo, hi in ((81, 20000, 32, 2500), (82, 1500, 0, 90)):
            off, data, meta, ts = _oracle_inputs(seed, n, lo, hi, 200, extra=[b"ha", b"a", b"A", b"xy"])
            exp, order = co.pipeline(d, off, data, meta, ts, rules_on=True, max_size=2500, nthreads=os.cpu_count() or 4)
            cases.append((_upload(torch, dev, off, data, meta, ts), exp, order))
        torch.cuda.synchronize()
        _run(eng, cases[0][0], static=True, use_ts=True)        # first call on this device: symbol lookup (and torch's own
        torch.cuda.synchronize()                                # lazy context work) happen before the measurement
        free_before = torch.cuda.mem_get_info()[0]
        for _ in range(2):
            for b, exp, order in cases:
                adm, res, got_order = _run(eng, b, static=True, use_ts=True)
                res = res.view(mq.RESULT_DTYPE).reshape(b["n"])
                for f in ("label", "reason", "sentiment", "question", "flags", "words", "queue_level", "score_rt", "score_hi",
                          "accepted", "admit"):
                    assert np.array_equal(res[f], exp[f]), f
                assert adm == len(order) and np.array_equal(got_order.astype(np.uint32), order)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free_before      # no cudaMalloc / cudaFree behind the calls
    finally:
        eng.close()


def test_static_scratch_submit_refuses_a_wide_dictionary():
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda:0")
    from tests import c5
    eng = mq.Engine(device=0)
    try:
        for p, ci, lit in c5.dictionary(3000):
            eng.add_literal(p, lit, ci)
        eng.commit()
        assert eng.dict_info()["states"] > 256
        b = _upload(torch, dev, *co.generate(91, 100, 16, 200))
        with pytest.raises(mq.TriageError) as ei:
            _run(eng, b, static=True, use_ts=False)
        assert ei.value.status == mq._capi.E_ARG and "padded-arena" in ei.value.message
    finally:
        eng.close()
