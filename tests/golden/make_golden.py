#!/usr/bin/env python
"""Generates the committed golden vectors under tests/golden/.

Run from the repo root:  python tests/golden/make_golden.py

Sources of truth, in order:
  * kat_reference.json — inputs verbatim from /root/reference/tests/preprocessor_test.go
    (K1..K7); expected values = the test's own assertions plus values derived by hand from
    internal/preprocessor/preprocessor.go.  Written literally below,
    NOT computed, so both oracles are checked against it.
  * edge_cases.json / c1_10k.npz — computed by the independent Python restatement
    (oracle/py_oracle.py); the C restatement and the CUDA path are tested against them.
The Go reference itself cannot be run in this image (no Go toolchain), so no fixture here
comes from executing the reference.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import c_oracle as co  # noqa: E402  (generator only)
from oracle import py_oracle as po  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

KATS = [
    # id, content, preset, meta(user_priority), expect
    dict(id="K1", ref="tests/preprocessor_test.go:27-41", content="This is an urgent request that needs attention",
         preset=3, user_priority=0, label=2, reason=3, words=8, sentiment=0, question=False, analyzed=True),
    dict(id="K2", ref="tests/preprocessor_test.go:48-64", content="Regular message content",
         preset=3, user_priority=2, label=2, reason=1, words=3, sentiment=0, question=False, analyzed=True),
    dict(id="K3", ref="tests/preprocessor_test.go:70-78", content="This is a regular message",
         preset=1, user_priority=0, label=1, reason=0, words=0, sentiment=0, question=False, analyzed=False),
    dict(id="K4", ref="tests/preprocessor_test.go:84-104", content="Regular message content",
         preset=3, user_priority=0, label=3, reason=0, words=3, sentiment=0, question=False, analyzed=True),
    dict(id="K5", ref="tests/preprocessor_test.go:110-123", content="I need this right now, it's immediate",
         preset=3, user_priority=0, label=1, reason=3, words=7, sentiment=0, question=False, analyzed=True),
    dict(id="K6", ref="tests/preprocessor_test.go:129-137", content="Can you help me with this problem?",
         preset=3, user_priority=0, label=3, reason=0, words=7, sentiment=0, question=True, analyzed=True),
    dict(id="K7", ref="tests/preprocessor_test.go:140-148", content="I'm very frustrated and angry with this service!",
         preset=3, user_priority=0, label=3, reason=0, words=8, sentiment=-1, question=False, analyzed=True),
]

E = []


def edge(content: bytes, preset=3, up=0, ud=0, mp=0, vip=0, note=""):
    E.append(dict(content_hex=content.hex(), preset=preset, user_priority=up, user_default=ud,
                  meta_priority=mp, vip=vip, note=note))


def build_edges():
    u = lambda s: s.encode("utf-8")  # noqa: E731
    # keywords: case folding, fold orbits, non-overlap, substring semantics
    for s in ["asap", "ASAP", "AsAp", "a\u017fap", "A\u017fAP", "\u017foon", "SOON", "monsoon", "soonsoon", "sooon",
              "asapasap", "asasap", "urgenturgent", "URGENT!", "right now", "RIGHT NOW", "Right  now", "right\u00a0now",
              "right\tnow", "rightnow", "immediateimmediate", "emergencyemergency", "\u0130mmediate", "\u0131mportant",
              "important", "IMPORTANT", "pr\u0130ority", "priority", "critical", "cr\u0130tical", "prioritycritical",
              "urgent asap", "asap urgent urgent", "soon soon immediate immediate", "emergency critical important",
              "i\u0307mmediate", "ur gent", "u\u200brgent", "\u212a", "ris\u212a", "as\u00e4p"]:
        edge(u(s), note="kw")
    # sentiment: whole-field equality after strings.ToLower
    for s in ["good", "Good", "GOOD", "good!", "goods", "good bad bad", "sat\u0130sfied", "SAT\u0130SF\u0130ED", "terr\u0130ble",
              "sat\u0131sfied", "happy happy angry", "\tgood\n", "good\u00a0bad", "goodbad", "good\u2003bad",
              "good\u0085bad bad x x x x x x x x x", "good\u1680bad\u2000bad",
              "good\u200agreat", "good\u2028great\u2029great\u202fbad\u205fbad", "good\u3000great", "good\u200bbad", "good\u180ebad", "good\x1cbad",
              "good\x1fbad", "good\x00bad", "good\x0bbad\x0cbad", "excellent", "EXCELLENT awful", "frustrated frustrated great",
              "angry", "awful\r\nawful", "  great  ", "great", "greater", "\u017fatisfied", "bad", "b\u00e4d", "BAD BAD good",
              "\u212a good", "good\ufeffbad", "good\u2007bad", "good\u200cbad"]:
        edge(u(s), note="sent")
    # question detection
    for s in ["what is", "WHAT is", "What  is", "somehow it", "who", "who ", "why?", "?", "x?y", "what\tis", "what\u00a0is",
              "wh\u0130o ", "where to", "nowhere fast", "when ", " when", "how", "HOW ARE YOU", "anyhow ", "why\n", "w ho ",
              "what\u3000", "?what", "wher e", "who\u200b ", "\uff1f", "what\x00 ", "W\u0127at "]:
        edge(u(s), note="q")
    # invalid / ragged UTF-8
    for b in [b"\xffgood\xff", b"good\xc2", b"\xc2 good", b"\xe2\x80good", b"\xe2\x80\x80good\xe2\x80\x80", b"a\xc5",
              b"a\xc5\xbfap", b"a\xc5\xbf\xbfap", b"\xc0\xa0", b"\xed\xa0\x80", b"\xf0\x9f\x98\x80", b"\xf4\x90\x80\x80",
              b"\xe2\x84", b"\xe2\x84\xaa", b"ris\xe2\x84", b"\xc4\xb0", b"sat\xc4sfied", b"sat\xc4\xb0sfied", b"\x80soon",
              b"so\x80on", b"\xe2soon", b"\xe2\x80\xa8good", b"\xe2\x80\xa7good", b"good\xe2\x80", b"good\xe2", b"\xe3\x80\x80",
              b"\xe3\x80", b"\xe1\x9a\x80x", b"\xe1\x9a\x81x", b"x\xc2\x85y", b"x\xc2\x86y", b"\xc2\xa0\xc2\xa0", b"\xe2\x81\x9fq",
              b"\xe2\x81\x9eq", b"\xef\xbf\xbd", b"\xe0\x80\xa0", b"\xf0\x80\x80\xa0", b"urgent\xc5", b"urgen\xc5t"]:
        edge(b, note="utf8")
    # word counts
    for s in ["", " ", "   ", "a", " a ", "a b", "a  b\t\tc\n\nd", "one,two three.four", "\u3000a\u3000b\u3000",
              "a\u00a0\u00a0b", "\u4e2d\u6587 \u6d88\u606f", "\u4e2d\u6587\u3000\u6d88\u606f", "a\x85b", "\u2000\u2001\u2002x\u2003\u2004y\u2005\u2006\u2007z\u2008\u2009\u200a"]:
        edge(u(s), note="words")
    # decision tree
    for preset in (0, 1, 2, 3, 4):
        edge(b"this is urgent", preset=preset, note="tree")
        edge(b"nothing here", preset=preset, note="tree")
    for up in (1, 2, 3, 4, 5):
        edge(b"this is urgent asap asap", preset=3, up=up, note="tree")
        edge(b"this is urgent", preset=0, up=up, ud=2, note="tree")
        edge(b"this is urgent", preset=4, up=up, note="tree")
    for ud in (1, 2, 3, 4):
        edge(b"immediate", preset=3, ud=ud, note="tree")
        edge(b"", preset=0, ud=ud, mp=1, note="tree")
    for mp in (0, 1, 2, 3, 4):
        edge(b"", preset=3, mp=mp, note="tree")
        edge(b"", preset=0, mp=mp, note="tree")
        edge(b"x", preset=0, mp=mp, note="tree")
    edge(b"", preset=3, up=5, mp=1, note="tree")
    edge(b"vip message", vip=1, note="tree")
    edge(b"z" * 10000, note="len10000")
    edge(b"z" * 10001, note="len10001")
    edge(b"urgent " + b"z" * 10001, vip=1, note="len>10000 vip")


def run_py(P, e):
    o = P.process(bytes.fromhex(e["content_hex"]), e["preset"], e["user_priority"], e["user_default"], e["meta_priority"])
    return dict(label=o["label"], reason=o["reason"], analyzed=o["analyzed"], tie=o["tie"], words=o["words"],
                sentiment=o["sentiment"], question=o["question"], has_analysis=o["has_analysis"],
                score_rt=o["scores"].get(1, 0), score_hi=o["scores"].get(2, 0))


def main():
    P = po.PyOracle()
    with open(os.path.join(HERE, "kat_reference.json"), "w") as f:
        json.dump(KATS, f, indent=1)
    build_edges()
    for e in E:
        e["expect"] = run_py(P, e)
    with open(os.path.join(HERE, "edge_cases.json"), "w") as f:
        json.dump(E, f, indent=0)

    # C1: 10K messages, len U[256,4096], default dictionary, queues cap 10 000 (config.go:157)
    seed, n = 0xC0FFEE01, 10000
    off, data, meta, ts = co.generate(seed, n, 256, 4096, edge=50)
    labels = np.zeros(n, np.uint8); reason = np.zeros(n, np.uint8); flags = np.zeros(n, np.uint8)
    words = np.zeros(n, np.uint32); sent = np.zeros(n, np.int8); ques = np.zeros(n, np.uint8)
    for i in range(n):
        c = bytes(data[int(off[i]):int(off[i + 1])])
        m = meta[i]
        o = P.process(c, int(m["preset"]), int(m["user_priority"]), int(m["user_default"]), int(m["meta_priority"]))
        labels[i] = o["label"]; reason[i] = o["reason"]; words[i] = o["words"]; sent[i] = o["sentiment"]
        ques[i] = o["question"]
        flags[i] = (1 if o["analyzed"] else 0) | (2 if o["tie"] else 0) | (4 if o["has_analysis"] else 0)
    lens = (off[1:] - off[:-1]).astype(np.int64)
    vip = (meta["flags"] & 1).astype(np.uint8)
    out = {}
    for mode, rules in (("A", False), ("B", True)):
        for cap in (10000, 300):
            final, acc, admit, level, order = po.admit_and_order(labels.tolist(), lens.tolist(), vip.tolist(),
                                                                 ts.tolist(), rules_on=rules, max_size=cap)
            k = f"{mode}_{cap}"
            out[f"final_{k}"] = np.array(final, np.uint8)
            out[f"accepted_{k}"] = np.array(acc, np.uint8)
            out[f"admit_{k}"] = np.array(admit, np.uint8)
            out[f"order_{k}"] = np.array(order, np.uint32)
    np.savez_compressed(os.path.join(HERE, "c1_10k.npz"), seed=np.uint64(seed), n=np.uint32(n),
                        arena_sha256=np.frombuffer(hashlib.sha256(data.tobytes()).digest(), np.uint8),
                        offsets_sha256=np.frombuffer(hashlib.sha256(off.tobytes()).digest(), np.uint8),
                        meta_sha256=np.frombuffer(hashlib.sha256(meta.tobytes()).digest(), np.uint8),
                        label=labels, reason=reason, flags=flags, words=words, sentiment=sent, question=ques, **out)
    print("wrote", len(KATS), "KATs,", len(E), "edge cases, C1 n =", n,
          "ties:", int(((flags & 2) != 0).sum()))


if __name__ == "__main__":
    main()
