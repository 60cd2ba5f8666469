"""CPU tests of the batch JSON ingest (api/handlers.go:160-176 ShouldBindJSON into models.Message) and of the analysis
emit (:180-190 json.Marshal(AnalyzeMessageContent)), engine-free through the host build of csrc/ingest.cc.  The
independent decoder is Python's json module; where encoding/json differs from it (case-insensitive field names, invalid
UTF-8 and lone surrogates -> U+FFFD, integers only for int fields, one value per body) the expected values are written
out by hand from the Go rules."""
import json

import numpy as np

from tests import hostcheck as hc


def _expect_meta(obj, users=None):
    """what ProcessMessage will look at, from a parsed JSON object with exact lower-case field names"""
    md = obj.get("metadata") or {}
    names = {"realtime": 1, "high": 2, "normal": 3, "low": 4}

    def code(v, unrec):
        if not isinstance(v, str):
            return 0
        return names.get(v.lower(), unrec)
    pr = obj.get("priority") or 0
    return dict(preset=255 if (pr < 0 or pr > 254) else pr, user_priority=code(md.get("user_priority"), 5), meta_priority=code(md.get("priority"), 0),
                flags=1 if "vip_user" in md else 0, user_default=(users or {}).get(obj.get("user_id") or "", 0))


def test_generated_messages_round_trip_through_json():
    rng = np.random.default_rng(5)
    objs = []
    for i in range(400):
        o = {"id": f"m{i}", "conversation_id": "c", "user_id": f"user{i % 7}", "content": "".join(
            chr(int(c)) for c in rng.choice([32, 34, 92, 47, 9, 10, 97, 98, 0xE9, 0x4E2D, 0x1F600, 0x17F, 0x212A, 1, 127], int(rng.integers(0, 60))))}
        if i % 3 == 0:
            o["priority"] = int(rng.integers(0, 6))
        if i % 4 == 0:
            o["metadata"] = {"user_priority": ["high", "LOW", "Realtime", "bogus", 5, None][i % 6], "priority": ["normal", "HİGH", 3][i % 3], "k": [1, {"a": None}]}
        if i % 5 == 0:
            o.setdefault("metadata", {})["vip_user"] = None
        if i % 11 == 0:
            o.update(status="pending", queue_name="", retry_count=1, max_retries=3, timeout=30000000000, created_at="2026-01-02T03:04:05.123Z",
                     scheduled_at=None, completed_at="2026-01-02T03:04:05+08:00")
        objs.append(o)
    bodies = [json.dumps(o, ensure_ascii=(i % 2 == 0)).encode() for i, o in enumerate(objs)]
    contents, meta, status = hc.ingest(bodies, user="user3", user_prio=2)
    assert not status.any()
    for o, c, m in zip(objs, contents, meta):
        assert c == o["content"].encode(), (o, c)
        want = _expect_meta(o, {"user3": 2})
        if isinstance((o.get("metadata") or {}).get("priority"), str) and "İ" in o["metadata"]["priority"]:
            want["meta_priority"] = 2                        # strings.ToLower(U+0130) = 'i' in Go ("HİGH" -> "high"), not in Python
        got = {k: int(m[k]) for k in want}
        assert got == want, (o, got, want)


def test_go_specific_decoding_rules():
    cases = [
        (b'{"Content":"a","PRIORITY":2,"User_ID":"u"}', b"a", dict(preset=2)),                        # field names under strings.EqualFold
        ('{"uſer_id":"user3","content":"x"}'.encode(), b"x", dict(user_default=2)),                # ... which folds U+017F onto s
        (b'{"content":"a","content":"b"}', b"b", {}),                                                   # a later duplicate overrides
        (b'{"content":"a\xffb\xe2\x80"}', "a�b��".encode(), {}),                       # invalid UTF-8 -> U+FFFD per byte
        (b'{"content":"\\ud83d\\ude00 \\ud83d x \\ude00"}', "\U0001F600 � x �".encode(), {}),   # surrogate pair / lone surrogates
        (b'{"content":null,"priority":null,"metadata":null}', b"", dict(preset=0)),                       # null leaves the zero value
        (b'null', b"", dict(preset=0)),
        (b'  {"content":"a"} trailing garbage', b"a", {}),                                               # Decoder.Decode reads one value
        (b'{"content":"a","metadata":{"vip_user":false,"user_priority":"high","user_priority":7}}', b"a", dict(flags=1, user_priority=0)),
        (b'{"content":"","metadata":{"priority":"Low"}}', b"", dict(meta_priority=4)),
        (b'{"content":"a","priority":-1}', b"a", dict(preset=255)), (b'{"content":"a","priority":300}', b"a", dict(preset=255)),
        (b'{"content":"a","unknown":{"deep":[1,2,{"x":"y"}]},"retry_count":0}', b"a", {}),
    ]
    bodies = [c[0] for c in cases]
    contents, meta, status = hc.ingest(bodies, user="user3", user_prio=2)
    for (body, want_c, want_m), c, m, st in zip(cases, contents, meta, status):
        assert st == 0 and c == want_c, (body, st, c)
        for k, v in want_m.items():
            assert int(m[k]) == v, (body, k, int(m[k]), v)


def test_what_the_handler_rejects_with_400():
    bad = [b'', b'{', b'{"content":"a"', b'[1,2]', b'"just a string"', b'42', b'{"content":5}', b'{"content":"a","priority":"3"}',
           b'{"content":"a","priority":3.0}', b'{"content":"a","priority":3e0}', b'{"content":"a","priority":99999999999999999999}',
           b'{"content":"a","retry_count":"x"}', b'{"content":"a","metadata":[1]}', b'{"content":"a","metadata":"m"}',
           b'{"content":"a","created_at":"yesterday"}', b'{"content":"a","created_at":"2026-02-30T00:00:00Z"}', b'{"content":"a","scheduled_at":12}',
           b'{"content":"a\nb"}', b'{"content":"a\\qb"}', b'{"content":"\\u12"}', b"{'content':'a'}", b'{"content":"a",}', b'{"content" "a"}',
           b'{"content":"a","timeout":true}', b'{"id":7}', b'{"content":"a","metadata":{"k":}}', b'{"content":tru}']
    good = b'{"content":"ok"}'
    bodies = []
    for b in bad:
        bodies += [b, good]
    contents, meta, status = hc.ingest(bodies)
    assert status[0::2].all() and not status[1::2].any()
    assert all(c == b"" for c in contents[0::2]) and all(c == b"ok" for c in contents[1::2])    # rejected bodies leave an empty slot


def test_analysis_json_is_what_encoding_json_writes():
    res = np.zeros(4, hc.RESULT_DTYPE)
    res["words"] = [0, 8, 123456, 7]
    res["sentiment"] = [0, -1, 1, 0]
    res["question"] = [0, 0, 1, 1]
    out = hc.emit(res)
    for r, o in zip(res, out):
        want = json.dumps({"word_count": int(r["words"]), "sentiment": {0: "neutral", 1: "positive", -1: "negative"}[int(r["sentiment"])],
                           "is_question": bool(r["question"])}, sort_keys=True, separators=(",", ":"))
        assert o.decode() == want       # Go sorts map keys; the values need no escaping, so there is no HTML-escaping difference
    assert out[1] == b'{"is_question":false,"sentiment":"negative","word_count":8}'
