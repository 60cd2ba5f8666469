"""Random dictionaries and matching text for differential tests (dictionary compiler + scan vs the oracle)."""
import numpy as np

ALPHABETS = [
    "abc", "abAB", "urgentasp", "xyz01", "aé批ſK", "no w", "abcdefghijklmnopqrstuvwxyz",
]


def fold(s: str) -> str:
    """regexp (?i) folding as far as it matters for borders: ASCII case, long s and the Kelvin sign"""
    return s.lower().replace("ſ", "s").replace("K", "k")


def bordered(lit: str, ci: bool) -> bool:
    t = fold(lit) if ci else lit
    b = t.encode()
    return any(b[:k] == b[-k:] for k in range(1, len(b)))


def make(seed: int, allow_bordered: bool, max_literals: int = 20):
    """returns (entries [(prio, ci, bytes)], contents [bytes])"""
    rng = np.random.default_rng(seed)
    alpha = ALPHABETS[int(rng.integers(len(ALPHABETS)))]
    entries, seen = [], set()
    for _ in range(int(rng.integers(1, max_literals + 1))):
        n = int(rng.integers(1, 11))
        lit = "".join(alpha[int(j)] for j in rng.integers(0, len(alpha), n))
        if lit.strip() != lit or not lit:       # no leading / trailing blanks: keeps the cases readable
            continue
        ci = bool(rng.integers(2)) and lit.isascii()   # "(?i)" literals with non-ASCII runes are refused by the API (and the oracle)
        if not allow_bordered and bordered(lit, ci):
            continue
        key = (fold(lit) if ci else lit, ci)
        if key in seen:
            continue
        seen.add(key)
        entries.append((int(rng.integers(1, 5)), ci, lit.encode()))
    if not entries:
        entries.append((2, True, b"qz"))
    fixed = ["good", "bad", "what", "who", "urgent", "asap", "great", "awful", "why"]
    contents = []
    for _ in range(int(rng.integers(60, 200))):
        parts = []
        for _ in range(int(rng.integers(0, 60))):
            r = rng.random()
            if r < 0.35:
                w = entries[int(rng.integers(len(entries)))][2].decode()
            elif r < 0.5:
                w = entries[int(rng.integers(len(entries)))][2].decode()
                w = w[: int(rng.integers(0, len(w) + 1))]
            elif r < 0.6:
                w = fixed[int(rng.integers(len(fixed)))]
            else:
                w = "".join(alpha[int(j)] for j in rng.integers(0, len(alpha), int(rng.integers(1, 8))))
            if rng.random() < 0.2:
                w = w.upper()
            elif rng.random() < 0.1:
                w = w.swapcase()
            parts.append(w)
            parts.append([" ", "", "  ", "\n", " ", "?"][int(rng.integers(6))] if rng.random() < 0.7 else "")
        contents.append("".join(parts).encode())
    return entries, contents
