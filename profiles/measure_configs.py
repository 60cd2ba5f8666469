#!/usr/bin/env python
"""Side measurements for DESIGN.md (not the bench contract): scan time of the other BASELINE configs on one B200.
  python profiles/measure_configs.py c3|c4|c5
c3: 1 Mi messages, mixed lengths (70 % 256-1024 B, 25 % 1-4 KB, 5 % 4-16 KB), default dictionary.
c4: 16 384 long-context payloads of 32-128 KB (quarter of SURVEY's 65 536), default dictionary.
c5: 262 144 messages of 256-4096 B, 10 000-keyword dictionary (wide table format)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import llm_message_queue_b200 as mq  # noqa: E402
from oracle import c_oracle as co  # noqa: E402
from tests import c5  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "c4"
e = mq.Engine(device=0)
if which == "c5":
    for p, ci, lit in c5.dictionary():
        e.add_literal(p, lit, ci)
    off, data, meta, ts = co.generate(0xC0FFEE05, 1 << 18, 256, 4096, mix=2, edge=50)
elif which == "c3":
    off, data, meta, ts = co.generate(0xC0FFEE03, 1 << 20, 256, 16384, mix=1, edge=50)
else:
    off, data, meta, ts = co.generate(0xC0FFEE04, 1 << 14, 32768, 131072, edge=50)
info = e.dict_info()
poff, plen, pbytes = e.pack_padded(off, data)
# device-resident submits (the host-buffer entry points pipeline the upload with the scan: their scan time includes waiting for chunks)
import torch  # noqa: E402
dev = torch.device("cuda", 0)
n = len(meta)
total = int(poff[-1])
d_bytes = torch.from_numpy(pbytes).to(dev)
d_len = torch.from_numpy(plen.view(np.int32)).to(dev)
d_off = torch.from_numpy(poff.view(np.int64)).to(dev)
d_meta = torch.from_numpy(meta.view(np.uint8).reshape(n, 16)).to(dev)
d_res = torch.zeros((n, 16), dtype=torch.uint8, device=dev)
d_ord = torch.zeros(n, dtype=torch.int32, device=dev)
times = []
for _ in range(8):
    e.submit_device_padded(d_off.data_ptr(), d_len.data_ptr(), d_bytes.data_ptr(), total, d_meta.data_ptr(), None, n, d_res.data_ptr(), d_ord.data_ptr())
    times.append(e.timings())
t = times[-1]
scan = min(x["scan_ms"] for x in times[2:])
print(json.dumps({"config": which, "messages": n, "arena_bytes": total, "automaton_states": info["states"],
                  "table_bytes": info["table_bytes"], "scan_variant": t["scan_variant"], "scan_row_bytes": t["scan_row_bytes"], "scan_ms": scan,
                  "scan_GBps": total / scan / 1e6, "finalize_ms": t["finalize_ms"], "order_ms": t["order_ms"],
                  "msgs_per_s_device": n / (t["total_ms"] / 1e3)}))
