"""GPU parity at the sizes that are timed (SURVEY 8d goldens): every output array of the CUDA path, for the whole batch,
against the SHA-256 digests the CPU oracle produced for the same seeded inputs (tests/golden/fullsize_sha256.json, made
by tests/golden/make_fullsize_digests.py).  Nothing is sampled: a single wrong label, count or position in the dequeue
order changes a digest.  Row kernel at its three row lengths, the mini-tile kernel, the packed-arena kernel and the wide
(10 000-keyword) kernel are all driven through the C ABI."""
import os

import numpy as np
import pytest

import llm_message_queue_b200 as mq
from tests import fullsize

pytestmark = pytest.mark.gpu


def _run_padded(e, poff, plen, pbytes, meta, ts, n):
    res, base = e.submit_padded(poff, plen, pbytes, meta, ts)
    handles, _ = e.drain(n)
    return res, (handles - base).astype(np.uint32)


def _compare(name, got, want, ctx):
    bad = [k for k in want if k in got and got[k] != want[k]]
    assert not bad, f"{name} ({ctx}): arrays that differ from the oracle's digests: {bad}"


@pytest.mark.parametrize("name", ["c2", "c3", "c4", "c4_rules"])
def test_default_dictionary_full_size_digests(name, monkeypatch):
    want = fullsize.load()[name]
    off, data, meta, ts, _, rules = fullsize.build(name)
    n = len(meta)
    assert n == want["messages"] and int(off[-1]) == want["content_bytes"]
    e = mq.Engine(device=0, rules_enabled=rules)
    try:
        poff, plen, pbytes = e.pack_padded(off, data)
        # default: the pair-lane kernel (two bytes per lookup); TRIAGE_PAIR=0: the one-lookup-per-byte row kernel
        variants = [("pair auto", {"TRIAGE_PAIR": "1"}), ("pair 1280", {"TRIAGE_PAIR": "1", "TRIAGE_ROW_BYTES": "1280"}),
                    ("pair 640", {"TRIAGE_PAIR": "1", "TRIAGE_ROW_BYTES": "640"}),
                    ("pair 320", {"TRIAGE_PAIR": "1", "TRIAGE_ROW_BYTES": "320"}), ("rows auto", {"TRIAGE_PAIR": "0"}), ("default", {}),
                    ("rows 1280", {"TRIAGE_PAIR": "0", "TRIAGE_ROW_BYTES": "1280"}), ("rows 320", {"TRIAGE_PAIR": "0", "TRIAGE_ROW_BYTES": "320"})]
        if name == "c2":
            variants += [("rows 640", {"TRIAGE_PAIR": "0", "TRIAGE_ROW_BYTES": "640"}), ("pair 20 warps", {"TRIAGE_PAIR": "1", "TRIAGE_PAIR_WARPS": "20"}),
                         ("mini-tile kernel", {"TRIAGE_ROWS": "0"})]
        for ctx, env in variants:
            for k in ("TRIAGE_ROW_BYTES", "TRIAGE_ROWS", "TRIAGE_PAIR", "TRIAGE_PAIR_WARPS"):
                monkeypatch.delenv(k, raising=False)
            for k, v in env.items():
                monkeypatch.setenv(k, v)
            if any(k in env for k in ("TRIAGE_ROWS", "TRIAGE_PAIR", "TRIAGE_PAIR_WARPS")):      # read at dictionary commit
                e2 = mq.Engine(device=0, rules_enabled=rules)
                try:
                    res, order = _run_padded(e2, poff, plen, pbytes, meta, ts, n)
                finally:
                    e2.close()
            else:
                res, order = _run_padded(e, poff, plen, pbytes, meta, ts, n)
            _compare(name, fullsize.digests(res, order), want, ctx)
        if name == "c2":                  # the packed wire format gives the same arrays
            for k in ("TRIAGE_ROW_BYTES", "TRIAGE_ROWS", "TRIAGE_PAIR", "TRIAGE_PAIR_WARPS"):
                monkeypatch.delenv(k, raising=False)
            res, base = e.submit(off, data, meta, ts)
            handles, _ = e.drain(n)
            _compare(name, fullsize.digests(res, (handles - base).astype(np.uint32)), want, "packed arena")
    finally:
        e.close()


def test_c5_10k_keywords_full_size_digests():
    want = fullsize.load()["c5"]
    off, data, meta, ts, entries, rules = fullsize.build("c5")
    n = len(meta)
    assert n == want["messages"] and int(off[-1]) == want["content_bytes"]
    e = mq.Engine(device=0, rules_enabled=rules)
    try:
        for p, ci, lit in entries:
            e.add_literal(p, lit, ci)
        e.commit()
        poff, plen, pbytes = e.pack_padded(off, data)
        res, order = _run_padded(e, poff, plen, pbytes, meta, ts, n)
        _compare("c5", fullsize.digests(res, order), want, "wide dictionary")
    finally:
        e.close()
