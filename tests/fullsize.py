"""The BASELINE configs at the sizes the benchmark times, shared by the digest generator
(tests/golden/make_fullsize_digests.py, runs the CPU oracle once, here) and the GPU tests that compare the CUDA path
with those digests array by array (tests/test_gpu_fullsize.py).  TEST INFRASTRUCTURE: uses oracle/ for the generator."""
import hashlib
import json
import os

import numpy as np

from oracle import c_oracle as co

HERE = os.path.dirname(os.path.abspath(__file__))
DIGESTS = os.path.join(HERE, "golden", "fullsize_sha256.json")

FIELDS = ("label", "reason", "sentiment", "question", "flags", "words", "queue_level", "score_rt", "score_hi", "accepted", "admit")

# name -> (seed, n, len_lo, len_hi, mix, rules_on, large dictionary)
CONFIGS = {
    "c2": (0xC0FFEE02, 1 << 20, 256, 4096, 0, False, False),          # configs[1]: 1 Mi messages, 256 B - 4 KB (what bench.py times)
    "c3": (0xC0FFEE03, 1 << 20, 256, 16384, 1, False, False),         # configs[2] length mix (70/25/5 %), one GPU's worth
    "c4": (0xC0FFEE04, 16384, 32768, 131072, 0, False, False),        # configs[3]: 32-128 KB payloads, rules off
    "c4_rules": (0xC0FFEE04, 16384, 32768, 131072, 0, True, False),   # ... rules on: every non-VIP message -> Low
    "c5": (0xC0FFEE05, 262144, 256, 4096, 2, False, True),            # configs[4]: skewed traffic, 10 000 keywords
}


def build(name):
    """-> (offsets, data, meta, ts, dictionary entries or None, rules_on)"""
    seed, n, lo, hi, mix, rules, large = CONFIGS[name]
    off, data, meta, ts = co.generate(seed, n, lo, hi, mix=mix, edge=50)
    entries = None
    if large:
        from tests import c5
        entries = c5.dictionary()
        contents = c5.inject([bytes(data[int(off[i]):int(off[i + 1])]) for i in range(n)], entries)
        off, data = co.pack(contents)
    return off, data, meta, ts, entries, rules


def digests(res, order):
    out = {f: hashlib.sha256(np.ascontiguousarray(res[f]).tobytes()).hexdigest() for f in FIELDS}
    out["order"] = hashlib.sha256(np.ascontiguousarray(order, dtype=np.uint32).tobytes()).hexdigest()
    out["n_admitted"] = int(len(order))
    return out


def load():
    with open(DIGESTS) as f:
        return json.load(f)
