"""Random patterns from the regex subset of AddKeywordPattern (tests only) and text that makes them fire."""
import numpy as np

ALPHA = "abcxyz"


def _atom(rng, depth):
    r = rng.integers(0, 100)
    if r < 45 or depth > 2:
        return ALPHA[int(rng.integers(0, len(ALPHA)))]
    if r < 55:
        return "."
    if r < 70:
        k = int(rng.integers(1, 4))
        body = "".join(ALPHA[int(j)] for j in rng.integers(0, len(ALPHA), k))
        if rng.integers(0, 4) == 0:
            body += "a-c"
        return "[" + ("^" if rng.integers(0, 5) == 0 else "") + body + "]"
    if r < 76:
        return ["\\d", "\\w", "\\s", "\\.", "\\D"][int(rng.integers(0, 5))]
    return "(" + ("?:" if rng.integers(0, 2) else "") + _alt(rng, depth + 1) + ")"


def _repeat(rng, depth):
    a = _atom(rng, depth)
    r = rng.integers(0, 100)
    if r < 60:
        return a
    q = ["?", "*", "+", "{2}", "{1,3}", "{2,}"][int(rng.integers(0, 6))]
    if rng.integers(0, 4) == 0:
        q += "?"
    return a + q


def _concat(rng, depth):
    return "".join(_repeat(rng, depth) for _ in range(int(rng.integers(1, 4))))


def _alt(rng, depth):
    return "|".join(_concat(rng, depth) for _ in range(1 if rng.integers(0, 3) else 2))


def pattern(rng):
    p = _alt(rng, 0)
    if rng.integers(0, 4) == 0:
        p = "(?i)" + p
    return p


def text(rng, n):
    """bytes over a small alphabet with spaces, capitals, digits, newlines and a few non-ASCII / invalid bytes"""
    pool = [c.encode() for c in ALPHA] * 6 + [c.upper().encode() for c in ALPHA] + [b" ", b" ", b"\n", b"1", b"7", b".", b"_",
                                                                                   "é".encode(), "ſ".encode(), "K".encode(), b"\xff", b"\xe2\x80"]
    return b"".join(pool[int(j)] for j in rng.integers(0, len(pool), n))
