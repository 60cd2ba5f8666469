"""GPU parity tests (run on the B200 box with -m gpu).  Every call goes through the C ABI of
libtriage_b200.so; the oracle (C restatement of the Go reference) and the committed goldens are the checkers.
Integer / byte work throughout: the bar is bit-exact, no tolerances."""
import json
import os

import numpy as np
import pytest

import llm_message_queue_b200 as mq
from oracle import c_oracle as co

pytestmark = pytest.mark.gpu

FIELDS = ("label", "reason", "sentiment", "question", "flags", "words", "queue_level", "score_rt", "score_hi")


@pytest.fixture(scope="module")
def eng():
    e = mq.Engine(device=0)
    yield e
    e.close()


def _assert_same(res, exp, fields=FIELDS, ctx=""):
    for f in fields:
        if not np.array_equal(res[f], exp[f]):
            bad = np.nonzero(res[f] != exp[f])[0]
            raise AssertionError(f"{ctx} field {f}: {len(bad)} mismatches, first {bad[:5]} got {res[f][bad[:5]]} want {exp[f][bad[:5]]}")


def _meta_from_cases(cases):
    meta = np.zeros(len(cases), dtype=mq.META_DTYPE)
    for i, e in enumerate(cases):
        meta[i]["preset"], meta[i]["user_priority"] = e["preset"], e["user_priority"]
        meta[i]["user_default"], meta[i]["meta_priority"], meta[i]["flags"] = e["user_default"], e["meta_priority"], e["vip"]
    return meta


def test_dictionary_facts(eng):
    info = eng.dict_info()
    assert 100 <= info["states"] <= 256 and info["warm_bytes"] == 16 and info["table_bytes"] == info["states"] * 516
    assert eng.dict_count(1) == 4 and eng.dict_count(2) == 5 and eng.dict_count(3) == 0


def test_edge_cases_vs_golden(eng, golden_dir):
    cases = json.load(open(os.path.join(golden_dir, "edge_cases.json")))
    contents = [bytes.fromhex(e["content_hex"]) for e in cases]
    off, data = co.pack(contents)
    res = eng.process(off, data, _meta_from_cases(cases))
    for i, e in enumerate(cases):
        x, r = e["expect"], res[i]
        got = dict(label=int(r["label"]), reason=int(r["reason"]), analyzed=bool(r["flags"] & 1), tie=bool(r["flags"] & 2),
                   words=int(r["words"]), sentiment=int(r["sentiment"]), question=bool(r["question"]),
                   has_analysis=bool(r["flags"] & 4), score_rt=int(r["score_rt"]), score_hi=int(r["score_hi"]))
        assert got == x, (i, e["note"], contents[i][:40], got, x)


@pytest.mark.parametrize("mode,rules", [("A", False), ("B", True)])
@pytest.mark.parametrize("cap", [10000, 300])
def test_c1_golden_labels_admission_and_dequeue_order(golden_dir, mode, rules, cap):
    g = np.load(os.path.join(golden_dir, "c1_10k.npz"))
    off, data, meta, ts = co.generate(int(g["seed"]), int(g["n"]), 256, 4096, edge=50)
    e = mq.Engine(device=0, rules_enabled=rules, max_queue_size=cap)
    try:
        res, base = e.submit(off, data, meta, ts)
        for key in ("reason", "flags", "words", "sentiment", "question"):
            assert np.array_equal(res[key], g[key]), key
        k = f"{mode}_{cap}"
        assert np.array_equal(res["label"], g[f"final_{k}"])
        assert np.array_equal(res["queue_level"], g["label"])          # level queue = pre-rule label
        assert np.array_equal(res["accepted"], g[f"accepted_{k}"])
        assert np.array_equal(res["admit"], g[f"admit_{k}"])
        handles, levels = e.drain(int(g["n"]) + 10)
        assert np.array_equal((handles - base).astype(np.uint32), g[f"order_{k}"])
        assert np.array_equal(levels, g["label"][g[f"order_{k}"]])
        assert e.drain(5)[0].size == 0
    finally:
        e.close()


@pytest.mark.parametrize("seed,n,lo,hi,edge,mix", [
    (11, 20000, 256, 4096, 50, 0),        # C2-shaped
    (12, 50000, 0, 48, 3000, 0),          # tiny + empty messages, heavy edge slice: many messages per tile
    (13, 300, 32768, 131072, 200, 0),     # C4: long-context payloads
    (14, 30000, 256, 16384, 50, 1),       # C3 length mix
    (15, 20000, 256, 4096, 10000, 0),     # every message from the edge generator
    (16, 3, 1, 5, 0, 0),
])
def test_generated_vs_c_oracle(eng, seed, n, lo, hi, edge, mix):
    off, data, meta, ts = co.generate(seed, n, lo, hi, mix=mix, edge=edge)
    exp, order = co.pipeline(co.Dict(), off, data, meta, ts, nthreads=os.cpu_count() or 4)
    res = eng.process(off, data, meta)
    _assert_same(res, exp, ctx=f"seed {seed}")
    e2 = mq.Engine(device=0)
    try:
        res2, base = e2.submit(off, data, meta, ts)
        _assert_same(res2, exp, FIELDS + ("accepted", "admit"), ctx=f"submit seed {seed}")
        handles, _ = e2.drain(n)
        assert np.array_equal((handles - base).astype(np.uint32), order)
    finally:
        e2.close()


def test_rules_and_capacity_vs_c_oracle():
    off, data, meta, ts = co.generate(21, 8000, 9000, 11000, edge=50)
    for rules, cap in ((True, 0), (True, 500), (False, 1)):
        exp, order = co.pipeline(co.Dict(), off, data, meta, ts, rules_on=rules, max_size=cap, nthreads=os.cpu_count() or 4)
        e = mq.Engine(device=0, rules_enabled=rules, max_queue_size=cap)
        try:
            res, base = e.submit(off, data, meta, ts)
            _assert_same(res, exp, FIELDS + ("accepted", "admit"), ctx=f"rules={rules} cap={cap}")
            handles, _ = e.drain(len(meta))
            assert np.array_equal((handles - base).astype(np.uint32), order)
        finally:
            e.close()


def test_non_monotone_enqueue_times_take_the_radix_path():
    n = 30000
    off, data, meta, _ = co.generate(31, n, 16, 300, edge=50)
    rng = np.random.default_rng(7)
    ts = rng.permutation(n).astype(np.uint64) * 1000 + np.uint64(1_700_000_000_000_000_000)   # unique, shuffled
    exp, order = co.pipeline(co.Dict(), off, data, meta, ts, nthreads=os.cpu_count() or 4)
    e = mq.Engine(device=0)
    try:
        res, base = e.submit(off, data, meta, ts)
        assert e.timings()["order_passes"] > 1
        handles, _ = e.drain(n)
        assert np.array_equal((handles - base).astype(np.uint32), order)
        # duplicates: container/heap is unstable on ties, the engine's convention is (priority, time, sequence)
        ts2 = (rng.integers(0, 50, n).astype(np.uint64) << np.uint64(33)) + np.uint64(5)
        res, base = e.submit(off, data, meta, ts2)
        handles, _ = e.drain(n)
        acc = np.nonzero(res["accepted"])[0]
        want = acc[np.lexsort((acc, ts2[acc], res["label"][acc]))]
        assert np.array_equal((handles - base).astype(np.int64), want)
    finally:
        e.close()


def test_two_submits_interleave_like_one_heap():
    off, data, meta, ts = co.generate(41, 6000, 64, 512, edge=50)
    n = len(meta)
    exp, order = co.pipeline(co.Dict(), off, data, meta, ts, max_size=700, nthreads=4)
    e = mq.Engine(device=0, max_queue_size=700)
    try:
        h = n // 3
        r1, b1 = e.submit(off[:h + 1], data, meta[:h], ts[:h])
        r2, b2 = e.submit(off[h:], data, meta[h:], ts[h:])      # offsets not starting at 0: rebased by the engine
        assert b2 == b1 + h
        res = np.concatenate([r1, r2])
        _assert_same(res, exp, FIELDS + ("accepted", "admit"))
        handles, _ = e.drain(n)
        assert np.array_equal((handles - b1).astype(np.uint32), order)
    finally:
        e.close()


def test_batch_into_one_named_queue_orders_by_priority_then_time():
    off, data, meta, ts = co.generate(51, 5000, 32, 256, edge=50)
    e = mq.Engine(device=0)
    try:
        q = e.queue_create(1200)
        res, base = e.submit(off, data, meta, ts, target_queue=q)
        lab = res["label"].astype(np.int64)
        assert res["accepted"].sum() == 1200 and (res["accepted"][:1200] == 1).all()   # capacity in arrival order
        got = np.array([e.pop(q) - base for _ in range(1200)])
        idx = np.arange(1200)
        assert np.array_equal(got, idx[np.argsort(lab[:1200], kind="stable")])
        with pytest.raises(mq.TriageError) as ei:
            e.pop(q)
        assert ei.value.status == mq._capi.E_EMPTY
    finally:
        e.close()


def test_full_size_c2_properties():
    """1M messages, 256 B - 4 KB (BASELINE configs[1]).  The oracle checks a sample; the whole batch is checked through
    size-independent properties: the order is the stable partition of admitted messages by label, every admitted message
    appears once, and per-label counts match."""
    n = 1 << 20
    off, data, meta, ts = co.generate(0xC0FFEE02, n, 256, 4096, edge=50)
    e = mq.Engine(device=0)
    try:
        res, base = e.submit(off, data, meta, None)
        t = e.timings()
        assert t["launches"] >= 5 and t["scan_ms"] > 0
        handles, levels = e.drain(n)
        order = (handles - base).astype(np.int64)
        acc = np.nonzero(res["accepted"])[0]
        assert len(order) == len(acc)
        assert np.array_equal(order, acc[np.argsort(res["label"][acc], kind="stable")])
        assert np.array_equal(levels, res["label"][order])
        rng = np.random.default_rng(1)
        sample = np.sort(rng.choice(n, 20000, replace=False))
        so, sd = co.pack([bytes(data[int(off[i]):int(off[i + 1])]) for i in sample])
        exp = co.process_batch(co.Dict(), so, sd, meta[sample], nthreads=os.cpu_count() or 4)
        _assert_same(res[sample], exp)
    finally:
        e.close()


def test_device_resident_entry_matches_host_entry(eng):
    torch = pytest.importorskip("torch")
    off, data, meta, ts = co.generate(61, 40000, 256, 2048, edge=50)
    n = len(meta)
    exp, order = co.pipeline(co.Dict(), off, data, meta, ts, nthreads=os.cpu_count() or 4)
    dev = torch.device("cuda:0")
    pad = np.zeros(((len(data) + 31) // 16) * 16, np.uint8)
    pad[:len(data)] = data
    d_bytes = torch.from_numpy(pad).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    d_meta = torch.from_numpy(meta.view(np.uint8).reshape(n, 16)).to(dev)
    d_ts = torch.from_numpy(ts.view(np.int64)).to(dev)
    d_res = torch.zeros((n, 16), dtype=torch.uint8, device=dev)
    d_ord = torch.zeros(n, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    for ts_ptr in (d_ts.data_ptr(), None):
        adm = eng.submit_device(d_off.data_ptr(), d_bytes.data_ptr(), len(data), d_meta.data_ptr(), ts_ptr, n,
                                d_res.data_ptr(), d_ord.data_ptr())
        res = d_res.cpu().numpy().view(mq.RESULT_DTYPE).reshape(n)
        _assert_same(res, exp, FIELDS + ("accepted", "admit"))
        assert adm == len(order)
        assert np.array_equal(d_ord.cpu().numpy()[:adm].astype(np.uint32), order)


@pytest.mark.parametrize("seed,n,lo,hi,edge,mix", [
    (111, 20000, 256, 4096, 50, 0),
    (112, 50000, 0, 48, 3000, 0),          # tiny + empty messages: more than 34 messages per mini-tile (global-offset list path)
    (113, 300, 32768, 131072, 200, 0),
    (114, 30000, 256, 16384, 50, 1),
    (115, 20000, 256, 4096, 10000, 0),
    (116, 3, 1, 5, 0, 0),
])
def test_padded_arena_path_vs_c_oracle(seed, n, lo, hi, edge, mix):
    """The engine's preferred wire format (16-byte aligned starts, reset-pad fill) through triage_pack_padded +
    triage_submit_padded: same bits as the oracle, rules and capacity included."""
    off, data, meta, ts = co.generate(seed, n, lo, hi, mix=mix, edge=edge)
    exp, order = co.pipeline(co.Dict(), off, data, meta, ts, rules_on=True, max_size=7000, nthreads=os.cpu_count() or 4)
    e = mq.Engine(device=0, rules_enabled=True, max_queue_size=7000)
    try:
        assert e.pad_byte() == 9
        poff, plen, pbytes = e.pack_padded(off, data)
        assert (poff % 16 == 0).all() and np.array_equal(plen, (off[1:] - off[:-1]).astype(np.uint32))
        res, base = e.submit_padded(poff, plen, pbytes, meta, ts)
        _assert_same(res, exp, FIELDS + ("accepted", "admit"), ctx=f"padded seed {seed}")
        handles, _ = e.drain(n)
        assert np.array_equal((handles - base).astype(np.uint32), order)
    finally:
        e.close()


@pytest.mark.parametrize("rows", ["1", "0", "1280", "640", "320", "p0-1", "p0-1280", "p0-320"])
def test_padded_arena_sizes_around_row_and_tile_boundaries(rows, monkeypatch):
    """The padded arena is scanned as 320-byte rows, 32 rows per tile (2-D TMA), with the partial last row read directly;
    arenas shorter than one row go to the mini-tile kernel.  Sweep totals across those boundaries, with both kernels
    (TRIAGE_ROWS=0 keeps to the mini-tile kernel): every result must equal the oracle's."""
    if rows.startswith("p0-"):            # the one-lookup-per-byte row kernel (the default) ...
        monkeypatch.setenv("TRIAGE_PAIR", "0")
        rows = rows[3:]
    else:                                 # ... and the pair-lane kernel
        monkeypatch.setenv("TRIAGE_PAIR", "1")
    if rows in ("1280", "640", "320"):   # force the long / short row length of the row-lane kernel whatever the arena size
        monkeypatch.setenv("TRIAGE_ROW_BYTES", rows)
    else:
        monkeypatch.setenv("TRIAGE_ROWS", rows)
    e = mq.Engine(device=0)
    try:
        seen = set()
        for seed, n, lo, hi in [(1, 1, 1, 15), (2, 1, 290, 303), (3, 2, 140, 160), (4, 3, 100, 110), (5, 40, 200, 300), (6, 33, 300, 304),
                                (7, 64, 150, 170), (8, 65, 304, 320), (9, 700, 0, 30), (10, 31, 305, 319), (11, 257, 1, 90), (12, 5, 2000, 2100),
                                (13, 900, 100, 1500), (14, 41, 1200, 1290), (15, 3000, 16, 400)]:
            off, data, meta, ts = co.generate(seed, n, lo, hi, edge=1500)
            exp, order = co.pipeline(co.Dict(), off, data, meta, ts, nthreads=4)
            poff, plen, pbytes = e.pack_padded(off, data)
            seen.add((int(poff[-1]) % 320) // 16)
            res, base = e.submit_padded(poff, plen, pbytes, meta, ts)
            _assert_same(res, exp, FIELDS, ctx=f"rows={rows} seed {seed} total {int(poff[-1])}")
            handles, _ = e.drain(n)
            assert np.array_equal((handles - base).astype(np.uint32), order)
        assert len(seen) >= 6   # several different tail-row lengths were hit
    finally:
        e.close()


@pytest.mark.parametrize("pair", ["1", "0"])
def test_event_dense_content_overflows_the_replay_queue(pair, monkeypatch):
    """Row-lane scans (pair lanes and one-lookup-per-byte lanes): events / flagged groups are queued per warp (24 slots) and
    replayed in bulk; content where almost every group carries an event overflows the queue, so lanes replay on the spot
    as well.  Both paths must add up to the oracle."""
    monkeypatch.setenv("TRIAGE_PAIR", pair)
    rng = np.random.default_rng(17)
    words = [b"asap", b"soon", b"urgent", b"good", b"bad", b"what", b"who", b"ASAP", "a\u017fap".encode(),
             b"x", b"critical", b"happy", b"why"]
    contents = []
    for i in range(600):
        k = int(rng.integers(1, 400))
        parts = [words[int(j)] for j in rng.integers(0, len(words), k)]
        sep = b" " if i % 3 else b"\n"
        contents.append(sep.join(parts) + (b"?" if i % 5 == 0 else b""))
    off, data = co.pack(contents)
    n = len(contents)
    _, _, meta, ts = co.generate(18, n, 1, 2)   # realistic meta records and enqueue times, own content
    exp, order = co.pipeline(co.Dict(), off, data, meta, ts, nthreads=4)
    e = mq.Engine(device=0)
    try:
        poff, plen, pbytes = e.pack_padded(off, data)
        res, base = e.submit_padded(poff, plen, pbytes, meta, ts)
        _assert_same(res, exp, FIELDS, ctx="event-dense")
        handles, _ = e.drain(n)
        assert np.array_equal((handles - base).astype(np.uint32), order)
    finally:
        e.close()


@pytest.mark.parametrize("seed", range(8))
def test_random_dictionaries_through_the_engine(seed):
    """differential: random literal dictionaries (self-overlapping ones included: fix-up kernel) added on top of the
    defaults, on text built from their own literals; labels, scores, analysis and dequeue order vs the oracle, through
    both wire formats"""
    from tests import randdict
    entries, contents = randdict.make(2000 + seed, allow_bordered=True)
    d = co.Dict()
    e = mq.Engine(device=0)
    try:
        for p, ci, lit in entries:
            assert d.add(p, lit, ci) == 0
            e.add_literal(p, lit, ci)
        off, data = co.pack(contents)
        n = len(contents)
        _, _, meta, ts = co.generate(3000 + seed, n, 1, 2)
        exp, order = co.pipeline(d, off, data, meta, ts, nthreads=4)
        e.commit()
        if e.dict_info()["states"] <= 256:   # larger automata use the wide table format: padded arena only
            res, base = e.submit(off, data, meta, ts)
            _assert_same(res, exp, FIELDS, ctx=f"random dictionary {seed} (packed) {entries}")
            handles, _ = e.drain(n)
            assert np.array_equal((handles - base).astype(np.uint32), order)
        if e.pad_byte() >= 0:
            poff, plen, pbytes = e.pack_padded(off, data)
            res, base = e.submit_padded(poff, plen, pbytes, meta, ts)
            _assert_same(res, exp, FIELDS, ctx=f"random dictionary {seed} (padded) {entries}")
            handles, _ = e.drain(n)
            assert np.array_equal((handles - base).astype(np.uint32), order)
    finally:
        e.close()


def test_prefetched_batches_give_the_same_results():
    """triage_prefetch uploads batch k+1 while batch k is processed; the matching submit (same host arrays) must return
    exactly what a plain submit returns — also when two batches are in flight, when a prefetch is never consumed, and
    for the packed arena."""
    batches = []
    for seed, n, lo, hi in [(301, 6000, 64, 2000), (302, 9000, 0, 300), (303, 500, 2000, 9000)]:
        off, data, meta, ts = co.generate(seed, n, lo, hi, edge=300)
        batches.append((off, data, meta, ts))
    ref, e = mq.Engine(device=0), mq.Engine(device=0)
    try:
        expected = []
        for off, data, meta, ts in batches:
            poff, plen, pbytes = ref.pack_padded(off, data)
            res, base = ref.submit_padded(poff, plen, pbytes, meta, ts)
            handles, _ = ref.drain(len(meta))
            expected.append((res, (handles - base).astype(np.uint32)))
        padded = [e.pack_padded(off, data) for off, data, _, _ in batches]
        # ring of two: prefetch k+1 before submitting k
        e.prefetch(padded[0][0], padded[0][1], padded[0][2], batches[0][2], batches[0][3])
        for k, (off, data, meta, ts) in enumerate(batches):
            if k + 1 < len(batches):
                nx = padded[k + 1]
                e.prefetch(nx[0], nx[1], nx[2], batches[k + 1][2], batches[k + 1][3])
            poff, plen, pbytes = padded[k]
            res, base = e.submit_padded(poff, plen, pbytes, meta, ts)
            assert e.timings()["h2d_ms"] < 1.0   # no upload inside the call
            handles, _ = e.drain(len(meta))
            for f in res.dtype.names:
                assert np.array_equal(res[f], expected[k][0][f]), (k, f)
            assert np.array_equal((handles - base).astype(np.uint32), expected[k][1])
        # a prefetch that is never consumed, then an unrelated plain submit, then the packed arena prefetched
        e.prefetch(padded[1][0], padded[1][1], padded[1][2], batches[1][2], batches[1][3])
        off, data, meta, ts = batches[2]
        res, base = e.submit(off, data, meta, ts)
        e.drain(len(meta))
        assert all(np.array_equal(res[f], expected[2][0][f]) for f in res.dtype.names)
        e.prefetch(off, None, data, meta, ts)
        res, base = e.submit(off, data, meta, ts)
        handles, _ = e.drain(len(meta))
        assert all(np.array_equal(res[f], expected[2][0][f]) for f in res.dtype.names)
        assert np.array_equal((handles - base).astype(np.uint32), expected[2][1])
    finally:
        ref.close()
        e.close()


def test_more_than_2_24_messages():
    """Message indices beyond 2^24 do not fit the 24-bit field of the row-lane scan's deferred-event queue: those lanes
    replay on the spot.  16.7 M empty messages followed by 3 000 keyword-dense ones: results and dequeue order of the
    tail (and the counts of the whole batch) against the oracle."""
    n_empty = (1 << 24) + 123
    off_t, data_t, meta_t, _ = co.generate(77, 3000, 16, 700, edge=2000)
    n = n_empty + len(meta_t)
    off = np.zeros(n + 1, dtype=np.uint64)
    off[n_empty:] = off_t
    meta = np.zeros(n, dtype=mq.META_DTYPE)
    meta["user_priority"] = 0
    meta[n_empty:] = meta_t
    ts = (np.arange(n, dtype=np.uint64) * np.uint64(7)) + np.uint64(10**15)
    exp, order = co.pipeline(co.Dict(), off, data_t, meta, ts, nthreads=os.cpu_count() or 4)
    e = mq.Engine(device=0)
    try:
        poff, plen, pbytes = e.pack_padded(off, data_t)
        assert int(poff[-1]) >= 16 * n_empty
        res, base = e.submit_padded(poff, plen, pbytes, meta, ts)
        _assert_same(res, exp, FIELDS + ("accepted", "admit"), ctx="n > 2^24")
        handles, _ = e.drain(n)
        assert np.array_equal((handles - base).astype(np.uint32), order)
    finally:
        e.close()


def test_padded_edge_cases_vs_golden(golden_dir):
    cases = json.load(open(os.path.join(golden_dir, "edge_cases.json")))
    contents = [bytes.fromhex(e["content_hex"]) for e in cases]
    off, data = co.pack(contents)
    e = mq.Engine(device=0)
    try:
        poff, plen, pbytes = e.pack_padded(off, data)
        res, _ = e.submit_padded(poff, plen, pbytes, _meta_from_cases(cases))
        for i, c in enumerate(cases):
            x, r = c["expect"], res[i]
            got = dict(label=int(r["label"]), reason=int(r["reason"]), analyzed=bool(r["flags"] & 1), tie=bool(r["flags"] & 2),
                       words=int(r["words"]), sentiment=int(r["sentiment"]), question=bool(r["question"]),
                       has_analysis=bool(r["flags"] & 4), score_rt=int(r["score_rt"]), score_hi=int(r["score_hi"]))
            assert got == x, (i, c["note"], contents[i][:40], got, x)
    finally:
        e.close()


def test_padded_full_size_c2_properties():
    n = 1 << 20
    off, data, meta, ts = co.generate(0xC0FFEE02, n, 256, 4096, edge=50)
    e = mq.Engine(device=0)
    try:
        poff, plen, pbytes = e.pack_padded(off, data)
        res, base = e.submit_padded(poff, plen, pbytes, meta, None)
        handles, levels = e.drain(n)
        order = (handles - base).astype(np.int64)
        acc = np.nonzero(res["accepted"])[0]
        assert np.array_equal(order, acc[np.argsort(res["label"][acc], kind="stable")])
        rng = np.random.default_rng(1)
        sample = np.sort(rng.choice(n, 20000, replace=False))
        so, sd = co.pack([bytes(data[int(off[i]):int(off[i + 1])]) for i in sample])
        exp = co.process_batch(co.Dict(), so, sd, meta[sample], nthreads=os.cpu_count() or 4)
        _assert_same(res[sample], exp)
        # and the packed path gives the very same records
        res2, base2 = e.submit(off, data, meta, None)
        assert res2.tobytes() == res.tobytes()
    finally:
        e.close()


def test_self_overlapping_keywords_count_without_overlap():
    """AddKeywordPattern("haha"): FindAllString counts leftmost NON-overlapping matches (preprocessor.go:146).  The chunked
    scan counts every occurrence and flags the message; the fix-up kernel recounts flagged messages sequentially."""
    extra = [(4, False, b"haha"), (3, False, b"abab"), (2, True, b"aa"), (1, False, b"xyxyx")]
    d = co.Dict()
    e = mq.Engine(device=0)
    try:
        for p, ci, lit in extra:
            assert d.add(p, lit, ci) == 0
            e.add_literal(p, lit, ci)
        rng = np.random.default_rng(9)
        off, data, meta, ts = co.generate(91, 4000, 16, 900, edge=100)
        contents = [bytes(data[int(off[i]):int(off[i + 1])]) for i in range(len(meta))]
        # runs long enough to straddle lanes, mini-tiles and CTAs
        for k in range(0, 4000, 7):
            unit = [b"ha", b"ab", b"a", b"A", b"xy", b"aA"][k % 6]
            reps = int(rng.integers(1, 4000 if k % 91 == 0 else 60))
            pos = int(rng.integers(0, len(contents[k]) + 1))
            contents[k] = contents[k][:pos] + unit * reps + contents[k][pos:]
        off, data = co.pack(contents)
        exp, order = co.pipeline(d, off, data, meta, ts, nthreads=os.cpu_count() or 4)
        assert (exp["score_hi"] > 3).sum() > 100      # the case is exercised
        res, base = e.submit(off, data, meta, ts)
        _assert_same(res, exp, FIELDS + ("accepted", "admit"), ctx="packed")
        e.drain(len(meta))
        poff, plen, pbytes = e.pack_padded(off, data)
        res, base = e.submit_padded(poff, plen, pbytes, meta, ts)
        _assert_same(res, exp, FIELDS + ("accepted", "admit"), ctx="padded")
        handles, _ = e.drain(len(meta))
        assert np.array_equal((handles - base).astype(np.uint32), order)
    finally:
        e.close()


def test_c5_large_dictionary_10k_keywords():
    """BASELINE configs[4]: 10 000 keywords added with AddKeywordPattern semantics (case-sensitive literals, ~1 %
    self-overlapping) on top of the defaults, skewed traffic.  The automaton (tens of thousands of states) takes the
    wide table format: class map in shared memory, transitions in L2-resident global memory."""
    from tests import c5
    entries = c5.dictionary()
    d = co.Dict()
    e = mq.Engine(device=0)
    try:
        for p, ci, lit in entries:
            assert d.add(p, lit, ci) == 0
            e.add_literal(p, lit, ci)
        info = e.dict_info()
        assert info["states"] > 50000
        off, data, meta, ts = co.generate(0xC0FFEE05, 3000, 256, 2048, mix=2, edge=50)
        contents = c5.inject([bytes(data[int(off[i]):int(off[i + 1])]) for i in range(len(meta))], entries)
        off, data = co.pack(contents)
        exp, order = co.pipeline(d, off, data, meta, ts, nthreads=os.cpu_count() or 4)
        assert len(set(exp["label"].tolist())) >= 4
        poff, plen, pbytes = e.pack_padded(off, data)
        res, base = e.submit_padded(poff, plen, pbytes, meta, ts)
        _assert_same(res, exp, FIELDS + ("accepted", "admit"), ctx="C5 padded")
        handles, _ = e.drain(len(meta))
        assert np.array_equal((handles - base).astype(np.uint32), order)
        res2 = e.process(off, data, meta)            # the mirror's ProcessMessage path re-packs for large dictionaries
        _assert_same(res2, exp, ctx="C5 process")
        with pytest.raises(mq.TriageError):          # the packed submit says so instead of silently doing something else
            e.submit(off, data, meta, ts)
    finally:
        e.close()
