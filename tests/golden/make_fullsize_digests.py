#!/usr/bin/env python
"""SHA-256 digests of every output array of the CPU oracle at the sizes the benchmark times ("Goldens for
C2-C5: SHA-256 over each output array, produced by the multithreaded CPU oracle").

    python tests/golden/make_fullsize_digests.py [name ...]      # from the repo root; minutes of CPU

Writes tests/golden/fullsize_sha256.json.  tests/test_gpu_fullsize.py recomputes the inputs from the same seeds on the
GPU box, runs the CUDA path and compares digest by digest, so the whole 1 Mi-message batch the row kernel is timed on
is compared with the oracle, not a sample.  The 10 000-keyword config uses the oracle's one-pass mode
(orc_dict_set_fast; checked against the per-pattern mode in tests/test_oracle.py) - 10 000 passes per message would take
days.  The Go reference cannot run here (no toolchain): like every golden in this directory these are oracle outputs.
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import c_oracle as co  # noqa: E402
from tests import fullsize  # noqa: E402


def main():
    names = sys.argv[1:] or list(fullsize.CONFIGS)
    out = {}
    if os.path.exists(fullsize.DIGESTS):
        out = fullsize.load()
    for name in names:
        t0 = time.time()
        off, data, meta, ts, entries, rules = fullsize.build(name)
        d = co.Dict()
        if entries is not None:
            for p, ci, lit in entries:
                assert d.add(p, lit, ci) == 0
            d.set_fast()
        res, order = co.pipeline(d, off, data, meta, ts, rules_on=rules, nthreads=os.cpu_count() or 4)
        rec = fullsize.digests(res, order)
        rec["messages"] = int(len(meta))
        rec["content_bytes"] = int(off[-1])
        rec["label_histogram"] = [int((res["label"] == p).sum()) for p in range(5)]
        rec["ties"] = int(((res["flags"] & 2) != 0).sum())
        out[name] = rec
        print(f"{name}: {len(meta)} messages, {int(off[-1])} bytes, {time.time() - t0:.0f} s, labels {rec['label_histogram']}", flush=True)
    with open(fullsize.DIGESTS, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
