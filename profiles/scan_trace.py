"""When does each warp of the pair-lane scan end?  Needs the measurement build of the library
(make -C llm-message-queue_b200/csrc OUT=$PWD/profiles/libtriage_trace.so EXTRA=-DTRIAGE_SCAN_TRACE_BUILD).
Runs the benchmark batch a few times device-resident, has the last scan write {start, end, tiles, SM} per warp and prints
how the end times and tile counts spread over warps and SMs.  Usage on a B200: python profiles/scan_trace.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import llm_message_queue_b200 as mq
mq._capi.LIB_PATH = os.path.join(ROOT, "profiles", "libtriage_trace.so")
from oracle import c_oracle as co

n = 1 << 20
dev = torch.device("cuda", 0)
off, data, meta, _ = co.generate(0xC0FFEE02, n, 256, 4096, edge=50)
eng = mq.Engine(device=0)
poff, plen, pbytes = eng.pack_padded(off, data)
total = int(poff[-1])
d_bytes = torch.from_numpy(pbytes).to(dev)
d_len = torch.from_numpy(plen.view(np.int32)).to(dev)
d_off = torch.from_numpy(poff.view(np.int64)).to(dev)
d_meta = torch.from_numpy(meta.view(np.uint8).reshape(n, 16)).to(dev)
d_res = torch.zeros((n, 16), dtype=torch.uint8, device=dev)
d_ord = torch.zeros(n, dtype=torch.int32, device=dev)
path = "/tmp/scan_trace.txt"
for i in range(6):
    if i == 5:
        os.environ["TRIAGE_SCAN_TRACE"] = path
    eng.submit_device_padded(d_off.data_ptr(), d_len.data_ptr(), d_bytes.data_ptr(), total, d_meta.data_ptr(), None, n,
                             d_res.data_ptr(), d_ord.data_ptr())
t = eng.timings()
rows = np.loadtxt(path, dtype=np.uint64)
w, t0, t1, tiles, sm = rows.T
worst_ns = tiles >> np.uint64(32)
tiles = tiles & np.uint64(0xFFFFFFFF)
worst_tile = sm >> np.uint64(32)
sm = sm & np.uint64(0xFFFFFFFF)
start = t0.min()
end = (t1 - start).astype(np.float64) / 1e3      # us
beg = (t0 - start).astype(np.float64) / 1e3
print(f"variant {t['scan_variant']} rows {t['scan_row_bytes']} scan {t['scan_ms']:.4f} ms (with the trace write); {len(w)} warps, {int(tiles.sum())} tiles")
print(f"warp start  us: min {beg.min():.1f} p50 {np.median(beg):.1f} max {beg.max():.1f}")
print(f"warp end    us: min {end.min():.1f} p10 {np.percentile(end, 10):.1f} p50 {np.median(end):.1f} p90 {np.percentile(end, 90):.1f} max {end.max():.1f}")
print(f"tiles per warp: min {tiles.min()} p50 {int(np.median(tiles))} max {tiles.max()}")
print(f"idle warp-time at the tail: {100 * (end.max() - end).mean() / end.max():.1f} % of the kernel")
per_sm_end = np.array([end[sm == s].max() for s in np.unique(sm)])
per_sm_tiles = np.array([tiles[sm == s].sum() for s in np.unique(sm)])
print(f"SM end      us: min {per_sm_end.min():.1f} p50 {np.median(per_sm_end):.1f} max {per_sm_end.max():.1f}; tiles per SM: min {per_sm_tiles.min()} max {per_sm_tiles.max()}")
dur = (t1 - t0).astype(np.float64) / 1e3 / np.maximum(tiles.astype(np.float64), 1)
print(f"us per tile by warp: min {dur.min():.2f} p50 {np.median(dur):.2f} max {dur.max():.2f}")
order = np.argsort(end)
print("last 8 warps to end (warp, SM, tiles, end us):", [(int(w[i]), int(sm[i]), int(tiles[i]), round(float(end[i]), 1)) for i in order[-8:]])
print("first 8 warps to end:", [(int(w[i]), int(sm[i]), int(tiles[i]), round(float(end[i]), 1)) for i in order[:8]])
slow = np.argsort(worst_ns)[::-1][:12]
row_bytes = t["scan_row_bytes"]
print("slowest tiles (tile, us, warp):", [(int(worst_tile[i]), round(float(worst_ns[i]) / 1e3, 1), int(w[i])) for i in slow])
print(f"slowest tile per warp, us: p50 {np.median(worst_ns) / 1e3:.1f} p90 {np.percentile(worst_ns, 90) / 1e3:.1f} max {worst_ns.max() / 1e3:.1f}")
for i in slow[:4]:
    lo = int(worst_tile[i]) * 32 * row_bytes
    seg = pbytes[lo:lo + 32 * row_bytes]
    hi_bytes = int((seg >= 0x80).sum())
    m0 = int(np.searchsorted(poff, lo, side="right")) - 1
    m1 = int(np.searchsorted(poff, lo + 32 * row_bytes, side="left"))
    print(f"  tile {int(worst_tile[i])}: bytes >= 0x80: {hi_bytes} of {len(seg)}; messages {m0}..{m1} ({m1 - m0}); sample: {bytes(seg[:120])!r}")
eng.close()
