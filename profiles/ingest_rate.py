"""Host throughput of the batch JSON ingest (ingest.cc, what triage_ingest_json runs; TRIAGE_INGEST_THREADS, default up to
16): request bodies of models.Message -> padded arena + meta.  Runs anywhere (no GPU): python profiles/ingest_rate.py"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import c_oracle as co
from tests import hostcheck as hc

n = 20000
off, data, meta, _ = co.generate(0xC0FFEE02, n, 256, 4096, edge=50)
raw = bytes(data)
bodies = []
for i in range(n):
    content = raw[int(off[i]):int(off[i + 1])].decode("utf-8", "replace")
    bodies.append(json.dumps({"id": f"m{i}", "user_id": f"u{i % 97}", "content": content, "priority": 0,
                              "metadata": {"user_priority": "high"} if i % 50 == 0 else {}}).encode())
boff = np.zeros(n + 1, np.uint64)
np.cumsum([len(b) for b in bodies], out=boff[1:])
blob = np.frombuffer(b"".join(bodies), np.uint8)
o = np.zeros(n + 1, np.uint64); lens = np.zeros(n, np.uint32); m = np.zeros(n, hc.META_DTYPE); st = np.zeros(n, np.uint8)
total = C.c_uint64(); err = C.create_string_buffer(256)
p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
lib = hc.lib()
lib.hostcheck_ingest(p(blob), p(boff), n, 9, None, 0, None, 0, p(o), p(lens), p(m), p(st), C.byref(total), err, 256)
arena = np.zeros(total.value, np.uint8)
best = 1e9
for _ in range(5):
    t0 = time.perf_counter()
    rc = lib.hostcheck_ingest(p(blob), p(boff), n, 9, None, 0, p(arena), len(arena), p(o), p(lens), p(m), p(st), C.byref(total), err, 256)
    best = min(best, time.perf_counter() - t0)
assert rc == 0 and int(st.sum()) == 0
t0 = time.perf_counter()
for b in bodies[:2000]:
    json.loads(b)
py = (time.perf_counter() - t0) * n / 2000
print(f"{n} bodies, {len(blob) / 1e6:.1f} MB of JSON -> {total.value / 1e6:.1f} MB of padded arena: {best * 1e3:.1f} ms on {os.environ.get('TRIAGE_INGEST_THREADS', 'min(cores, 16)')} thread(s), {os.cpu_count()} cores = "
      f"{len(blob) / best / 1e6:.0f} MB/s = {n / best / 1e6:.2f} M messages/s (CPython json.loads on the same bodies: {py * 1e3:.0f} ms)")
