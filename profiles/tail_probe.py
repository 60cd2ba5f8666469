"""Where does the scan's tail come from?  Times the scan kernel (engine timings, CUDA events) on the C2 batch as the benchmark generates it (0.5 % of the
messages are edge cases: multi-byte whitespace and fold runes everywhere, every group goes through the exact walk; the
rest carries near-misses of the dictionary) and on plain words only.  Usage on a B200: TRIAGE_PAIR=0|1 python profiles/tail_probe.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_message_queue_b200 as mq
if os.environ.get("TRIAGE_LIB"):          # A/B of two builds on one box
    mq._capi.LIB_PATH = os.environ["TRIAGE_LIB"]
from oracle import c_oracle as co

n = 1 << 20
dev = torch.device("cuda", 0)
# (edge=0 means "the generator's default", 50 per 10 000 messages; only plain=True leaves the edge-case messages out)
for name, kw in (("benchmark set", dict(edge=50)), ("plain words", dict(edge=0, plain=True))):
    off, data, meta, _ = co.generate(0xC0FFEE02, n, 256, 4096, **kw)
    eng = mq.Engine(device=0)
    poff, plen, pbytes = eng.pack_padded(off, data)
    total = int(poff[-1])
    d_bytes = torch.from_numpy(pbytes).to(dev)
    d_len = torch.from_numpy(plen.view(np.int32)).to(dev)
    d_off = torch.from_numpy(poff.view(np.int64)).to(dev)
    d_meta = torch.from_numpy(meta.view(np.uint8).reshape(n, 16)).to(dev)
    d_res = torch.zeros((n, 16), dtype=torch.uint8, device=dev)
    d_ord = torch.zeros(n, dtype=torch.int32, device=dev)
    ms = []
    for i in range(13):
        eng.submit_device_padded(d_off.data_ptr(), d_len.data_ptr(), d_bytes.data_ptr(), total, d_meta.data_ptr(), None, n,
                                 d_res.data_ptr(), d_ord.data_ptr())
        t = eng.timings()
        if i >= 3:
            ms.append(t["scan_ms"])
    print(f"{name:18s} variant {t['scan_variant']} rows {t['scan_row_bytes']} arena {total / 1e9:.3f} GB scan {sum(ms) / len(ms):.4f} ms "
          f"= {total / (sum(ms) / len(ms)) / 1e6:.0f} GB/s", flush=True)
    eng.close()
    del d_bytes
