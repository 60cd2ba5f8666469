"""GPU test of the wire-format row: request bodies -> triage_ingest_json -> triage_submit_padded ->
triage_emit_analysis_json, against the oracle fed with the same messages decoded by Python's json module
(api/handlers.go:160-202: ShouldBindJSON, ProcessMessage, AnalyzeMessageContent + json.Marshal, PushMessage)."""
import json
import os

import numpy as np
import pytest

import llm_message_queue_b200 as mq
from oracle import c_oracle as co

pytestmark = pytest.mark.gpu


def test_json_bodies_through_the_engine_and_back():
    off, data, meta, ts = co.generate(0x1501, 5000, 0, 900, edge=300)
    n = len(meta)
    names = {1: "realtime", 2: "High", 3: "normal", 4: "LOW", 5: "whatever"}
    users = {"u-default-high": 2, "u-default-low": 4}
    bodies, contents = [], []
    for i in range(n):
        raw = bytes(data[int(off[i]):int(off[i + 1])])
        text = raw.decode("utf-8", errors="replace")          # JSON carries text: invalid bytes become U+FFFD on the way in
        contents.append(text.encode())
        o = {"id": f"m{i}", "user_id": "u-default-high" if meta[i]["user_default"] == 2 else ("u-default-low" if meta[i]["user_default"] == 4 else f"user{i}"),
             "content": text, "priority": int(meta[i]["preset"])}
        md = {}
        if meta[i]["user_priority"]:
            md["user_priority"] = names[int(meta[i]["user_priority"])]
        if meta[i]["meta_priority"]:
            md["priority"] = names[int(meta[i]["meta_priority"])]
        if meta[i]["flags"] & 1:
            md["vip_user"] = True
        if md or i % 2:
            o["metadata"] = md
        bodies.append(json.dumps(o, ensure_ascii=bool(i % 2)).encode())
    bodies[17] = b'{"content": 5}'                               # the handler answers 400: skipped
    bodies[18] = b'{"content":"x","priority":"high"}'
    # the oracle's view: same contents, meta as ProcessMessage will see it
    ometa = meta.copy()
    ometa["user_default"] = np.where(np.isin(meta["user_default"], (2, 4)), meta["user_default"], 0)
    ooff, odata = co.pack(contents)
    e = mq.Engine(device=0)
    try:
        for u, p in users.items():
            e.set_user_priority(u, p)
        poff, plen, arena, gmeta, status = e.ingest_json(bodies)
        assert status.sum() == 2 and status[17] and status[18]
        keep = np.nonzero(status == 0)[0]
        for f in ("preset", "user_priority", "user_default", "meta_priority"):
            assert np.array_equal(gmeta[f][keep], ometa[f][keep]), f
        assert np.array_equal(gmeta["flags"][keep] & 1, ometa["flags"][keep] & 1) and (gmeta["flags"] & mq._capi.META_ANALYZE).all()
        assert np.array_equal(plen[keep], (ooff[1:] - ooff[:-1]).astype(np.uint32)[keep])
        res, base = e.submit_padded(poff, plen, arena, gmeta, ts)
        exp, order = co.pipeline(co.Dict(), ooff, odata, ometa, ts, nthreads=os.cpu_count() or 4)
        for f in ("label", "reason", "flags", "queue_level", "score_rt", "score_hi"):
            assert np.array_equal(res[f][keep], exp[f][keep]), f
        analysed = keep[(exp["flags"][keep] & 4) != 0]
        for f in ("words", "sentiment", "question"):
            assert np.array_equal(res[f][analysed], exp[f][analysed]), f
        # Metadata["analysis"] (handlers.go:180-190): AnalyzeMessageContent also for messages ProcessMessage left alone
        out = e.emit_analysis_json(res)
        for i in keep[:1500]:
            w, p, q, isq = co.analyze(contents[i])
            want = json.dumps({"word_count": w, "sentiment": "positive" if p > q else ("negative" if q > p else "neutral"), "is_question": isq},
                              sort_keys=True, separators=(",", ":"))
            assert out[i].decode() == want, (i, contents[i][:50], out[i], want)
        e.drain(n)
    finally:
        e.close()
