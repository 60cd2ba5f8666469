"""How much of the scan time is bank conflicts?  Same kernel, same sizes, texts with different (state, byte) diversity."""
import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
import llm_message_queue_b200 as mq
n, mlen = 262144, 2176
rng = np.random.default_rng(3)
def build(kind):
    if kind == "all_a": body = np.full(n * mlen, ord('a'), np.uint8)
    elif kind == "two_letters": body = rng.choice(np.frombuffer(b"ab", np.uint8), n * mlen)
    elif kind == "az_uniform": body = rng.choice(np.frombuffer(b"abcdefghijklmnopqrstuvwxyz", np.uint8), n * mlen)
    elif kind == "az_space": body = rng.choice(np.frombuffer(b"abcdefghijklmnopqrstuvwxyz     ", np.uint8), n * mlen)
    elif kind == "random_bytes": body = rng.integers(32, 127, n * mlen).astype(np.uint8)
    off = (np.arange(n + 1, dtype=np.uint64) * np.uint64(mlen))
    return off, body
dev = torch.device("cuda", 0)
eng = mq.Engine(device=0)
for kind in ["all_a", "two_letters", "az_uniform", "az_space", "random_bytes"]:
    off, data = build(kind)
    meta = np.zeros(n, mq.META_DTYPE) if hasattr(mq, "META_DTYPE") else np.zeros((n, 16), np.uint8)
    off, lens, pad = eng.pack_padded(off, data)
    total = int(off[-1])
    d_bytes = torch.from_numpy(pad).to(dev); d_len = torch.from_numpy(lens.view(np.int32)).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev); d_meta = torch.from_numpy(meta.view(np.uint8).reshape(n, 16)).to(dev)
    d_res = torch.zeros((n, 16), dtype=torch.uint8, device=dev); d_ord = torch.zeros(n, dtype=torch.int32, device=dev)
    ts = []
    for i in range(8):
        eng.submit_device_padded(d_off.data_ptr(), d_len.data_ptr(), d_bytes.data_ptr(), total, d_meta.data_ptr(), None, n, d_res.data_ptr(), d_ord.data_ptr())
        t = eng.timings()
        if i >= 3: ts.append(t["scan_ms"])
    print(f"{kind:14s} bytes {total/1e6:7.1f} MB  scan {np.mean(ts):.4f} ms  {total/np.mean(ts)/1e6:8.1f} GB/s", flush=True)
