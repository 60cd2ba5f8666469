"""CPU tests of the oracle itself: C restatement vs the reference KATs, the committed goldens
(made by the independent Python restatement) and the container/heap KATs.
Reference tests mirrored: tests/preprocessor_test.go, tests/priorityqueue_test.go:27-97."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import c_oracle as co
from oracle import py_oracle as po


def _meta(preset=3, up=0, ud=0, mp=0, vip=0):
    m = np.zeros(1, dtype=co.META_DTYPE)
    m["preset"], m["user_priority"], m["user_default"], m["meta_priority"], m["flags"] = preset, up, ud, mp, vip
    return m


def _one(d, content: bytes, **kw):
    off, data = co.pack([content])
    return co.process_batch(d, off, data, _meta(**kw))[0]


def test_reference_kats_c_and_py(golden_dir):
    kats = json.load(open(os.path.join(golden_dir, "kat_reference.json")))
    d = co.Dict()
    P = po.PyOracle()
    assert len(kats) == 7
    for k in kats:
        c = k["content"].encode()
        r = _one(d, c, preset=k["preset"], up=k["user_priority"])
        o = P.process(c, k["preset"], k["user_priority"])
        for got in (dict(label=int(r["label"]), reason=int(r["reason"]), analyzed=bool(r["flags"] & 1),
                         words=int(r["words"]), sentiment=int(r["sentiment"]), question=bool(r["question"])),
                    {x: o[x] for x in ("label", "reason", "analyzed", "words", "sentiment", "question")}):
            for key, v in got.items():
                assert v == k[key], (k["id"], key, v, k[key])


def test_edge_cases_c_vs_golden(golden_dir):
    cases = json.load(open(os.path.join(golden_dir, "edge_cases.json")))
    d = co.Dict()
    contents = [bytes.fromhex(e["content_hex"]) for e in cases]
    off, data = co.pack(contents)
    meta = np.zeros(len(cases), dtype=co.META_DTYPE)
    for i, e in enumerate(cases):
        meta[i]["preset"], meta[i]["user_priority"] = e["preset"], e["user_priority"]
        meta[i]["user_default"], meta[i]["meta_priority"], meta[i]["flags"] = e["user_default"], e["meta_priority"], e["vip"]
    res = co.process_batch(d, off, data, meta, nthreads=3)
    for i, e in enumerate(cases):
        x, r = e["expect"], res[i]
        got = dict(label=int(r["label"]), reason=int(r["reason"]), analyzed=bool(r["flags"] & 1), tie=bool(r["flags"] & 2),
                   words=int(r["words"]), sentiment=int(r["sentiment"]), question=bool(r["question"]),
                   has_analysis=bool(r["flags"] & 4), score_rt=int(r["score_rt"]), score_hi=int(r["score_hi"]))
        assert got == x, (i, e["note"], contents[i][:40], got, x)


def test_c1_generator_is_deterministic_and_matches_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "c1_10k.npz"))
    off, data, meta, ts = co.generate(int(g["seed"]), int(g["n"]), 256, 4096, edge=50)
    assert hashlib.sha256(data.tobytes()).digest() == g["arena_sha256"].tobytes()
    assert hashlib.sha256(off.tobytes()).digest() == g["offsets_sha256"].tobytes()
    assert hashlib.sha256(meta.tobytes()).digest() == g["meta_sha256"].tobytes()
    d = co.Dict()
    res = co.process_batch(d, off, data, meta, nthreads=4)
    for key in ("label", "reason", "flags", "words", "sentiment", "question"):
        assert np.array_equal(res[key], g[key]), key
    for mode, rules in (("A", False), ("B", True)):
        for cap in (10000, 300):
            r = res.copy()
            order = co.admit_and_order(r, meta, off, ts, rules_on=rules, max_size=cap)
            k = f"{mode}_{cap}"
            assert np.array_equal(r["label"], g[f"final_{k}"]), k
            assert np.array_equal(r["accepted"], g[f"accepted_{k}"]), k
            assert np.array_equal(r["admit"], g[f"admit_{k}"]), k
            assert np.array_equal(order, g[f"order_{k}"]), k


def test_order_is_stable_partition_by_label_mode_a(golden_dir):
    """Size-independent property the engine relies on: with strictly increasing
    timestamps the heap drain equals a stable sort by label."""
    off, data, meta, ts = co.generate(77, 5000, 64, 512, edge=50)
    res, order = co.pipeline(co.Dict(), off, data, meta, ts, nthreads=2)
    lab = res["label"].astype(np.int64)
    idx = np.nonzero(res["accepted"])[0]
    expect = idx[np.argsort(lab[idx], kind="stable")]
    assert np.array_equal(order, expect.astype(np.uint32))


def test_heap_matches_python_restatement_with_equal_timestamps():
    rng = np.random.default_rng(5)
    h = co.lib().orc_heap_new()
    g = po.GoHeap()
    import ctypes as C
    n = 2000
    pr = rng.integers(1, 5, n)
    ts = rng.integers(0, 50, n)  # many ties: exercises container/heap mechanics, not just Less
    for i in range(n):
        co.lib().orc_heap_push(h, int(pr[i]), int(ts[i]), i)
        g.push(int(pr[i]), int(ts[i]), i)
    out = C.c_uint32()
    for _ in range(n):
        assert co.lib().orc_heap_pop(h, C.byref(out)) == 1
        assert out.value == g.pop()
    assert co.lib().orc_heap_pop(h, C.byref(out)) == 0  # ErrQueueEmpty (tests/priorityqueue_test.go:94-97)
    co.lib().orc_heap_free(h)


def test_threads_do_not_change_results():
    off, data, meta, ts = co.generate(123, 3000, 16, 700, edge=300)
    d = co.Dict()
    a = co.process_batch(d, off, data, meta, nthreads=1)
    b = co.process_batch(d, off, data, meta, nthreads=7)
    assert a.tobytes() == b.tobytes()


def test_custom_patterns_case_sensitive_and_bordered():
    # AddKeywordPattern compiles the raw pattern: case-sensitive unless it carries (?i) (preprocessor.go:176-184)
    d = co.Dict()
    assert d.add(4, b"haha", False) == 0
    assert d.add(3, b"abab", False) == 0
    assert d.add(1, "紧急".encode(), False) == 0
    assert d.add(2, b"\xff", False) == 4
    assert d.add(2, "é".encode(), True) == 4
    P = po.PyOracle()
    P.add(4, "haha", False); P.add(3, "abab", False); P.add(1, "紧急", False)
    for c in [b"hahahaha", b"hahaha", b"HAHA haha", b"abababab", b"ababab", "紧急 紧急".encode(), b"x\xe7\xb4\xa7\xe6\x80x"]:
        sc = co.scores(d, c)
        pc = P.scores(po.decode(c))
        assert sc[1:] == [pc.get(p, 0) for p in (1, 2, 3, 4)], c
    assert co.scores(d, b"hahahaha")[4] == 2 and co.scores(d, b"hahaha")[4] == 1


def test_fast_mode_of_the_oracle_gives_the_per_pattern_scores():
    """orc_dict_set_fast (one trie pass for border-free case-sensitive literals, byte search for self-overlapping ones) is
    what produced the full-size digests of the 10 000-keyword config; it must equal the per-pattern passes that follow
    preprocessor.go:144-149 literally - checked here on traffic that carries the dictionary, self-overlaps included."""
    from tests import c5
    entries = c5.dictionary(2500)
    slow, fast = co.Dict(), co.Dict()
    for p, ci, lit in entries:
        assert slow.add(p, lit, ci) == 0 and fast.add(p, lit, ci) == 0
    for lit in (b"na\xc3\xafve", "étéété".encode(), b"zz"):   # non-ASCII literals, bordered ones
        assert slow.add(3, lit, False) == 0 and fast.add(3, lit, False) == 0
    fast.set_fast()
    off, data, meta, ts = co.generate(0xFA57, 400, 64, 1500, mix=2, edge=400)
    contents = c5.inject([bytes(data[int(off[i]):int(off[i + 1])]) for i in range(len(meta))], entries)
    contents[3] += "naïve étéétéété zzzzz \xc3".encode("latin-1", "ignore") + b"\xc3\xa9t\xc3" + "été".encode()
    off, data = co.pack(contents)
    a = co.process_batch(slow, off, data, meta, nthreads=4)
    b = co.process_batch(fast, off, data, meta, nthreads=4)
    assert a.tobytes() == b.tobytes()
    assert (a["label"] == 4).sum() > 50 and len(set(a["label"].tolist())) >= 3


def test_full_size_digests_are_committed_for_every_config():
    from tests import fullsize
    dig = fullsize.load()
    for name, cfg in fullsize.CONFIGS.items():
        assert name in dig and dig[name]["messages"] == cfg[1]
        for f in fullsize.FIELDS + ("order",):
            assert len(dig[name][f]) == 64
    assert dig["c4_rules"]["label_histogram"][4] > 16000      # rules on: every non-VIP long message -> Low


def test_next_backoff_three_ways():
    """ExponentialBackoff.NextBackoff (worker.go:274-290): the engine's C-ABI function (pure arithmetic, no GPU), the C
    oracle and the formula written out here agree bit for bit, cap and truncation included."""
    import llm_message_queue_b200 as mq
    for initial, mx, factor in ((10**9, 60 * 10**9, 2.0), (10**8, 10**10, 1.5), (250 * 10**6, 3 * 10**9, 1.1), (1, 2**62, 3.7), (5 * 10**9, 10**9, 2.0)):
        for rc in range(-1, 40):
            if rc <= 0:
                want = initial
            else:
                b = float(initial)
                for _ in range(1, rc):
                    b *= factor
                want = mx if b > float(mx) else int(b)
            assert co.next_backoff_ns(initial, mx, factor, rc) == want
            assert mq._capi.backoff_ns(initial, mx, factor, rc) == want
