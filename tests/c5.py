"""BASELINE configs[4] / SURVEY 8(d) C5: a 10 000-keyword dictionary (len 4-16, case-sensitive as AddKeywordPattern gives,
8000 Low / 1000 Normal / 700 High / 300 Realtime, ~1 % deliberately self-overlapping) and skewed traffic that carries it."""
import numpy as np

BORDERED = [b"haha", b"abab", b"xyzxyz", b"momo", b"lolol", b"tktk", b"nono", b"ququ", b"bubub", b"zaza"]


def dictionary(n=10000, seed=0xC5):
    rng = np.random.default_rng(seed)
    prios = [4] * 8000 + [3] * 1000 + [2] * 700 + [1] * 300
    rng.shuffle(prios)
    seen, out = set(), []
    while len(out) < n:
        if len(out) % 100 == 7:   # ~1 %: a proper prefix equals a suffix
            base = BORDERED[(len(out) // 100) % len(BORDERED)]
            k = base + bytes([97 + (len(out) // 1000)]) * 0 + bytes(rng.integers(97, 123, int(rng.integers(0, 2))).astype(np.uint8))
            k = k + k[:2] if len(k) % 2 else k
        else:
            k = bytes(rng.integers(97, 123, int(rng.integers(4, 17))).astype(np.uint8))
        if k in seen:
            continue
        seen.add(k)
        out.append((prios[len(out) % len(prios)], False, k))
    return out


def inject(contents, entries, seed=1, rate=6):
    """sprinkles dictionary keywords (sometimes repeated back to back, so self-overlaps happen) into messages"""
    rng = np.random.default_rng(seed)
    out = []
    for c in contents:
        for _ in range(int(rng.integers(0, rate))):
            k = entries[int(rng.integers(0, len(entries)))][2]
            pos = int(rng.integers(0, len(c) + 1))
            c = c[:pos] + k * int(rng.integers(1, 4)) + c[pos:]
        out.append(c)
    return out
