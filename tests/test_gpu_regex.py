"""GPU tests of AddKeywordPattern with regular expressions (preprocessor.go:176-184 compiles any pattern; :146 counts
len(FindAllString)): the RE2 subset the engine compiles to a rune-level program and walks on the device, against the C
oracle's backtracking matcher - labels, scores, analysis and dequeue order, both wire formats, the static entry too."""
import os

import numpy as np
import pytest

import llm_message_queue_b200 as mq
from oracle import c_oracle as co
from tests import randregex

pytestmark = pytest.mark.gpu

FIELDS = ("label", "reason", "sentiment", "question", "flags", "words", "queue_level", "score_rt", "score_hi", "accepted", "admit")


def _check(e, d, off, data, meta, ts, ctx):
    n = len(meta)
    exp, order = co.pipeline(d, off, data, meta, ts, nthreads=os.cpu_count() or 4)
    if e.dict_info()["states"] <= 256:
        res, base = e.submit(off, data, meta, ts)
        for f in FIELDS:
            assert np.array_equal(res[f], exp[f]), (ctx, "packed", f, np.nonzero(res[f] != exp[f])[0][:5])
        handles, _ = e.drain(n)
        assert np.array_equal((handles - base).astype(np.uint32), order), (ctx, "packed order")
    poff, plen, pbytes = e.pack_padded(off, data)
    res, base = e.submit_padded(poff, plen, pbytes, meta, ts)
    for f in FIELDS:
        assert np.array_equal(res[f], exp[f]), (ctx, "padded", f, np.nonzero(res[f] != exp[f])[0][:5])
    handles, _ = e.drain(n)
    assert np.array_equal((handles - base).astype(np.uint32), order), (ctx, "padded order")
    return exp


def test_regex_patterns_on_generated_traffic():
    pats = [(1, b"(?i)p[0-9]+ (incident|outage)s?"), (2, b"ur+gent(ly)?"), (4, b"\\w+@\\w+\\.(com|org)"), (3, b"[0-9]{3}-[0-9]{4}"),
            (4, b"(?i)unsub(scribe)?"), (2, b"(foo|foobar)baz"), (1, b"a{4,}?h")]
    d = co.Dict()
    e = mq.Engine(device=0)
    try:
        for p, rx in pats:
            assert d.add_regex(p, rx) == 0
            e.add_pattern(p, rx)
        assert e.dict_count(1) == 4 + 2 and e.dict_count(4) == 2
        off, data, meta, ts = co.generate(0x4E47, 6000, 16, 1500, edge=300)
        rng = np.random.default_rng(5)
        inserts = [b"P1 incident", b"p42 outages", b"urrrrgently", b"urgent", b"bob@example.com", b"a_b@x.org", b"555-1234", b"UNSUB", b"unsubscribe",
                   b"foobarbaz", b"foobaz", b"aaaaaaah", b"aaah", "urgént".encode(), b"p9 incident\xff", b"12-3456 123-45678"]
        contents = [bytes(data[int(off[i]):int(off[i + 1])]) for i in range(len(meta))]
        for k in range(0, len(contents), 3):
            for _ in range(int(rng.integers(1, 4))):
                ins = inserts[int(rng.integers(0, len(inserts)))]
                pos = int(rng.integers(0, len(contents[k]) + 1))
                contents[k] = contents[k][:pos] + b" " + ins + b" " + contents[k][pos:]
        off, data = co.pack(contents)
        exp = _check(e, d, off, data, meta, ts, "generated")
        assert len(set(exp["label"].tolist())) >= 4 and (exp["score_rt"] > 0).sum() > 100
    finally:
        e.close()


@pytest.mark.parametrize("seed", range(4))
def test_random_regex_dictionaries_vs_oracle(seed):
    rng = np.random.default_rng(7000 + seed)
    d = co.Dict()
    e = mq.Engine(device=0)
    try:
        added = 0
        while added < 6:
            pat = randregex.pattern(rng).encode()
            prio = int(rng.integers(1, 5))
            try:
                e.add_pattern(prio, pat)
            except mq.TriageError as err:
                assert err.status == mq._capi.E_PATTERN
                continue
            assert d.add_regex(prio, pat) == 0 or d.add(prio, pat.replace(b"\\", b""), False) == 0, pat   # a plain literal goes the literal way
            added += 1
        n = 1500
        contents = [randregex.text(rng, int(rng.integers(0, 400))) for _ in range(n)]
        off, data = co.pack(contents)
        _, _, meta, ts = co.generate(7100 + seed, n, 1, 2)
        _check(e, d, off, data, meta, ts, f"random regex dictionary {seed}")
    finally:
        e.close()
