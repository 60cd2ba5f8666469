"""py_oracle.py — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Second, independent CPU restatement (Python) of llm-message-queue's
preprocess -> priority-order path.  It exists only to cross-check the C
restatement (oracle/triage_oracle.c) and to generate the committed golden
vectors under tests/golden/.  Only tests/ and tests/golden/make_golden.py import it.

Parity pin: the Go reference cannot be built here; pinned against the KATs of
/root/reference/tests/preprocessor_test.go (K1..K7) and cross-checked with the C
restatement.  Multi-message dequeue order: parity unpinned by any reference test.

Where the C oracle walks runes by hand, this one leans on Python's `re` engine
with *explicit* character classes (never re.IGNORECASE, whose case folding is
wider than Go's unicode.SimpleFold: Python folds U+0130/U+0131 onto i, Go does
not), so the two restatements share no code and little structure.

Reference lines followed (relative to /root/reference):
  internal/preprocessor/preprocessor.go:24-53, 56-114, 117-168, 197-249
  internal/priorityqueue/queue.go:22-27, 92-139
  internal/priorityqueue/queue_manager.go:451-466
  internal/priorityqueue/queue_factory.go:211-233
  cmd/queue-manager/main.go:111-124
"""
from __future__ import annotations

import re
from dataclasses import dataclass, field

REALTIME, HIGH, NORMAL, LOW = 1, 2, 3, 4
NAMES = {1: "realtime", 2: "high", 3: "normal", 4: "low"}  # pkg/models/message.go:24-37
REASONS = {0: None, 1: "user_override", 2: "user_default", 3: "content_keywords"}

# unicode.IsSpace (Go): note U+001C..U+001F are NOT space (Python's str.split disagrees)
GO_SPACE = ("\t\n\v\f\r \x85\xa0\u1680" + "".join(chr(c) for c in range(0x2000, 0x200B))
            + "\u2028\u2029\u202f\u205f\u3000")
_FIELD_RE = re.compile("[^" + re.escape(GO_SPACE) + "]+")

# strings.ToLower as far as an all-ASCII needle can see it: A-Z, U+0130 -> i, U+212A -> k
_LOWER = {c: c + 32 for c in range(ord("A"), ord("Z") + 1)}
_LOWER[0x130] = ord("i")
_LOWER[0x212A] = ord("k")

POSITIVE = ["good", "great", "excellent", "happy", "satisfied"]  # preprocessor.go:49
NEGATIVE = ["bad", "terrible", "awful", "angry", "frustrated"]  # :50
QUESTION = ["what", "how", "why", "when", "where", "who"]  # :51


def decode(content: bytes) -> str:
    """Go ranges over a string rune by rune; an invalid byte is one RuneError of width 1.
    surrogateescape gives the same segmentation: each invalid byte becomes one lone surrogate."""
    return content.decode("utf-8", errors="surrogateescape")


def _ci_class(ch: str) -> str:
    """regexp "(?i)" literal: the unicode.SimpleFold orbit of an ASCII letter."""
    lo, up = ch.lower(), ch.upper()
    cls = {lo, up}
    if lo == "s":
        cls.add("\u017f")
    if lo == "k":
        cls.add("\u212a")
    return "[" + "".join(re.escape(c) for c in sorted(cls)) + "]"


@dataclass
class Pattern:
    prio: int
    literal: str
    ci: bool
    rx: re.Pattern = field(init=False)

    def __post_init__(self):
        parts = []
        for ch in self.literal:
            if self.ci and ch.isascii() and ch.isalpha():
                parts.append(_ci_class(ch))
            else:
                parts.append(re.escape(ch))
        self.rx = re.compile("".join(parts))


_PERL = {"d": "0-9", "w": "0-9A-Za-z_", "s": "\\t\\n\\f\\r "}   # RE2: ASCII only (Python's own \\d \\w \\s are Unicode-wide)


def _fold_members(lo: int, hi: int) -> str:
    """what "(?i)" adds to the ASCII range lo..hi: the other case, plus U+212A for k and U+017F for s"""
    out = []
    for c in range(lo, hi + 1):
        ch = chr(c)
        if ch.isascii() and ch.isalpha():
            out.append(ch.swapcase())
            if ch.lower() == "k":
                out.append("\u212a")
            if ch.lower() == "s":
                out.append("\u017f")
    return "".join(re.escape(x) for x in out)


def go_regex_to_py(pattern: str) -> "re.Pattern":
    """Go (RE2) pattern of the subset the engine accepts -> an equivalent Python `re` pattern with every class written out
    (no re.IGNORECASE, no re.ASCII): a token-level rewrite, Python's backtracking engine then supplies the
    leftmost-first semantics.  Raises ValueError for what this rewrite does not cover."""
    fold = dotall = False
    i = 0
    m = re.match(r"\(\?([is]+)\)", pattern)
    if m:
        fold, dotall, i = "i" in m.group(1), "s" in m.group(1), m.end()
    out = []
    n = len(pattern)

    def escape_item(j):
        """after a backslash at pattern[j]: -> (text for inside a class or None, code point or None, next index)"""
        c = pattern[j]
        if c in _PERL:
            return _PERL[c], None, j + 1
        if c in "DWS":
            return None, None, j + 1
        if c in "ntrfv":
            return None, ord({"n": "\n", "t": "\t", "r": "\r", "f": "\f", "v": "\v"}[c]), j + 1
        if c == "x":
            return None, int(pattern[j + 1:j + 3], 16), j + 3
        if not c.isalnum():
            return None, ord(c), j + 1
        raise ValueError("escape outside the subset")

    while i < n:
        c = pattern[i]
        if c == "\\":
            txt, cp, j = escape_item(i + 1)
            if txt is not None:
                out.append("[" + txt + "]")
            elif cp is None:
                out.append("[^" + _PERL[pattern[i + 1].lower()] + "]")
            else:
                out.append("[" + re.escape(chr(cp)) + (_fold_members(cp, cp) if fold and cp < 128 else "") + "]")
            i = j
        elif c == "[":
            i += 1
            neg = i < n and pattern[i] == "^"
            if neg:
                i += 1
            body, first = [], True
            while True:
                if pattern[i] == "]" and not first:
                    i += 1
                    break
                first = False
                if pattern[i] == "\\":
                    txt, cp, i = escape_item(i + 1)
                    if txt is not None:
                        body.append(txt)
                        continue
                    if cp is None:
                        raise ValueError("negated Perl class inside a class: not covered by the rewrite")
                    lo = cp
                else:
                    lo = ord(pattern[i]); i += 1
                hi = lo
                if pattern[i] == "-" and pattern[i + 1] != "]":
                    i += 1
                    if pattern[i] == "\\":
                        _, hi, i = escape_item(i + 1)
                    else:
                        hi = ord(pattern[i]); i += 1
                body.append(re.escape(chr(lo)) + ("-" + re.escape(chr(hi)) if hi != lo else ""))
                if fold:
                    if hi >= 128:
                        raise ValueError("non-ASCII under (?i)")
                    body.append(_fold_members(lo, hi))
            out.append("[" + ("^" if neg else "") + "".join(body) + "]")
        elif c == ".":
            out.append("[\\s\\S]" if dotall else "[^\\n]")
            i += 1
        elif c in "()|?*+{},^$" or c.isdigit() and False:
            out.append(c)
            i += 1
        else:
            if fold and ord(c) >= 128:
                raise ValueError("non-ASCII under (?i)")
            out.append("[" + re.escape(c) + _fold_members(ord(c), ord(c)) + "]" if fold and c.isascii() and c.isalpha() else re.escape(c))
            i += 1
    return re.compile("".join(out))


class PyOracle:
    def __init__(self, with_defaults: bool = True, default_prio: int = NORMAL):
        self.patterns: list[Pattern] = []
        self.default_prio = default_prio
        if with_defaults:  # preprocessor.go:28-43
            for w in ["immediate", "emergency", "asap", "right now"]:
                self.add(REALTIME, w, True)
            for w in ["urgent", "important", "priority", "critical", "soon"]:
                self.add(HIGH, w, True)

    def add(self, prio: int, literal: str, ci: bool) -> None:
        self.patterns.append(Pattern(prio, literal, ci))

    def add_regex(self, prio: int, pattern: str) -> None:
        """AddKeywordPattern(prio, pattern) with a regular expression (preprocessor.go:176-184)"""
        p = Pattern(prio, "x", False)
        p.rx = go_regex_to_py(pattern)
        self.patterns.append(p)

    # -- a5 --------------------------------------------------------------
    def scores(self, s: str) -> dict[int, int]:
        sc: dict[int, int] = {}
        for p in self.patterns:  # FindAllString(content, -1): leftmost, non-overlapping
            sc[p.prio] = sc.get(p.prio, 0) + sum(1 for _ in p.rx.finditer(s))
        return sc

    # -- a6 --------------------------------------------------------------
    @staticmethod
    def analyze(content: bytes):
        s = decode(content)
        fields = _FIELD_RE.findall(s)
        pos = neg = 0
        for f in fields:
            fl = f.translate(_LOWER)
            pos += sum(1 for w in POSITIVE if fl == w)
            neg += sum(1 for w in NEGATIVE if fl == w)
        q = content.endswith(b"?")
        if not q:
            low = s.translate(_LOWER)
            q = any((k + " ") in low for k in QUESTION)
        sentiment = 1 if pos > neg else (-1 if neg > pos else 0)
        return len(fields), sentiment, bool(q)

    # -- a4 --------------------------------------------------------------
    def process(self, content: bytes, preset=NORMAL, user_priority=0, user_default=0, meta_priority=0):
        """returns dict(label, reason, analyzed, tie, words, sentiment, question, has_analysis)"""
        out = dict(label=preset, reason=0, analyzed=False, tie=False, words=0, sentiment=0,
                   question=False, has_analysis=False, scores={})
        if preset not in (0, NORMAL):  # :63-65
            return out
        label, reason, tie = preset, 0, False
        if user_priority != 0:  # :68-82 (5 = unrecognised string: branch taken, nothing set)
            if 1 <= user_priority <= 4:
                label, reason = user_priority, 1
        elif user_default != 0:  # :83-86
            label, reason = user_default, 2
        else:
            if len(content) == 0:  # :120-138
                label = meta_priority if 1 <= meta_priority <= 4 else self.default_prio
            else:
                sc = self.scores(decode(content))
                out["scores"] = sc
                best = max(sc.values(), default=0)
                if best > 0:
                    winners = sorted(p for p, v in sc.items() if v == best)
                    label = winners[0]  # convention: most urgent wins; reference is random here
                    tie = len(winners) > 1
                else:
                    label = self.default_prio
            if label != preset:
                reason = 3
        out.update(label=label, reason=reason, analyzed=True, tie=tie)
        if len(content) > 0:  # :199-201
            w, s, q = self.analyze(content)
            out.update(words=w, sentiment=s, question=q, has_analysis=True)
        return out


# -- container/heap, restated -------------------------------------------------
class GoHeap:
    """container/heap over QueueItem with Less = (Priority, Timestamp) — queue.go:22-27."""

    def __init__(self):
        self.a: list[tuple[int, int, int]] = []

    @staticmethod
    def _less(x, y):
        if x[0] == y[0]:
            return x[1] < y[1]
        return x[0] < y[0]

    def push(self, prio, ts, ident):
        a = self.a
        a.append((prio, ts, ident))
        j = len(a) - 1
        while True:  # up
            i = (j - 1) // 2
            if i == j or j == 0 or not self._less(a[j], a[i]):
                break
            a[i], a[j] = a[j], a[i]
            j = i

    def pop(self):
        a = self.a
        n = len(a) - 1
        a[0], a[n] = a[n], a[0]
        i = 0
        while True:  # down(0, n)
            j1 = 2 * i + 1
            if j1 >= n or j1 < 0:
                break
            j = j1
            if j1 + 1 < n and self._less(a[j1 + 1], a[j1]):
                j = j1 + 1
            if not self._less(a[j], a[i]):
                break
            a[i], a[j] = a[j], a[i]
            i = j
        return a.pop()[2]

    def __len__(self):
        return len(self.a)


def admit_and_order(labels, lens, vip, ts, rules_on=False, max_size=0):
    """a9-a12.  Returns (final_labels, accepted, admit_code, queue_level, order)."""
    queues = {p: GoHeap() for p in (1, 2, 3, 4)}
    final = list(labels)
    accepted = [0] * len(labels)
    admit = [0] * len(labels)
    level = list(labels)
    for i, lab in enumerate(labels):
        if lab not in queues:
            admit[i] = 1  # ErrQueueNotFound: only the four level queues exist
            continue
        if rules_on:  # queue_manager.go:451-466 / queue_factory.go:211-233
            if vip[i]:
                final[i] = HIGH
            elif lens[i] > 10000:
                final[i] = LOW
        q = queues[lab]
        if max_size > 0 and len(q) >= max_size:  # queue.go:101-103
            admit[i] = 2
            continue
        accepted[i] = 1
        q.push(final[i], ts[i], i)
    order = []
    for p in (1, 2, 3, 4):  # cmd/queue-manager/main.go:112
        q = queues[p]
        while len(q):
            order.append(q.pop())
    return final, accepted, admit, level, order
