"""c_oracle.py — TEST INFRASTRUCTURE: ctypes binding of oracle/liboracle.so.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this.  The product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

META_DTYPE = np.dtype([("preset", "u1"), ("user_priority", "u1"), ("user_default", "u1"),
                       ("meta_priority", "u1"), ("flags", "u1"), ("reserved", "u1", (11,))])
RESULT_DTYPE = np.dtype([("label", "u1"), ("reason", "u1"), ("sentiment", "i1"), ("question", "u1"),
                         ("accepted", "u1"), ("admit", "u1"), ("flags", "u1"), ("queue_level", "u1"),
                         ("words", "<u4"), ("score_rt", "<u2"), ("score_hi", "<u2")])
assert META_DTYPE.itemsize == 16 and RESULT_DTYPE.itemsize == 16

RF_ANALYZED, RF_TIE, RF_HAS_ANALYSIS = 1, 2, 4


class GenCfg(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n", C.c_uint32), ("len_lo", C.c_uint32), ("len_hi", C.c_uint32),
                ("mix", C.c_uint32), ("edge_permille_x10", C.c_uint32), ("plain", C.c_uint32)]


def build(force: bool = False) -> str:
    src = [os.path.join(_HERE, f) for f in ("triage_oracle.c", "regex_oracle.c", "gen.c", "triage_oracle.h")]
    if force or not os.path.exists(_LIB_PATH) or any(os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        vp, u64p = C.c_void_p, C.c_void_p
        L.orc_dict_new.restype = vp
        L.orc_dict_new.argtypes = [C.c_int]
        L.orc_dict_add.restype = C.c_int
        L.orc_dict_add.argtypes = [vp, C.c_int, C.c_char_p, C.c_uint32, C.c_int]
        L.orc_dict_free.argtypes = [vp]
        L.orc_dict_set_fast.argtypes = [vp, C.c_int]
        L.orc_dict_add_regex.restype = C.c_int
        L.orc_dict_add_regex.argtypes = [vp, C.c_int, C.c_char_p, C.c_uint32]
        L.orc_regex_compile.restype = vp
        L.orc_regex_compile.argtypes = [C.c_char_p, C.c_uint32]
        L.orc_regex_free.argtypes = [vp]
        L.orc_regex_count.restype = C.c_uint32
        L.orc_regex_count.argtypes = [vp, C.c_char_p, C.c_uint64]
        L.orc_scores.argtypes = [vp, C.c_char_p, C.c_uint64, C.POINTER(C.c_uint32)]
        L.orc_analyze.argtypes = [C.c_char_p, C.c_uint64, C.POINTER(C.c_uint32), C.POINTER(C.c_int32),
                                  C.POINTER(C.c_int32), C.POINTER(C.c_int)]
        L.orc_process_batch.argtypes = [vp, C.c_int, u64p, vp, vp, C.c_uint32, vp, C.c_int]
        L.orc_admit_and_order.restype = C.c_uint32
        L.orc_admit_and_order.argtypes = [vp, vp, u64p, u64p, C.c_uint32, C.c_int, C.c_int64, vp]
        L.orc_pipeline.restype = C.c_uint32
        L.orc_pipeline.argtypes = [vp, C.c_int, u64p, vp, vp, u64p, C.c_uint32, C.c_int, C.c_int64, vp, vp, C.c_int]
        L.orc_generate.restype = C.c_uint64
        L.orc_generate.argtypes = [C.POINTER(GenCfg), u64p, vp, vp, u64p]
        L.orc_next_backoff_ns.restype = C.c_int64
        L.orc_next_backoff_ns.argtypes = [C.c_int64, C.c_int64, C.c_double, C.c_int]
        L.orc_heap_new.restype = vp
        L.orc_heap_free.argtypes = [vp]
        L.orc_heap_push.argtypes = [vp, C.c_int, C.c_uint64, C.c_uint32]
        L.orc_heap_pop.restype = C.c_int
        L.orc_heap_pop.argtypes = [vp, C.POINTER(C.c_uint32)]
        L.orc_heap_len.restype = C.c_uint32
        L.orc_heap_len.argtypes = [vp]
        _lib = L
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Dict:
    def __init__(self, with_defaults=True):
        self.h = lib().orc_dict_new(1 if with_defaults else 0)

    def add(self, prio: int, literal: bytes, ci: bool) -> int:
        return lib().orc_dict_add(self.h, prio, literal, len(literal), 1 if ci else 0)

    def add_regex(self, prio: int, pattern: bytes) -> int:
        """AddKeywordPattern(prio, pattern) with a regular expression (RE2 subset); 0 ok, 4 outside the subset"""
        return lib().orc_dict_add_regex(self.h, prio, pattern, len(pattern))

    def set_fast(self, on=True):
        """one-pass scoring for border-free case-sensitive literals (same numbers; for the full-size 10 K-keyword goldens).
        Call after the last add()."""
        lib().orc_dict_set_fast(self.h, 1 if on else 0)

    def __del__(self):
        try:
            lib().orc_dict_free(self.h)
        except Exception:
            pass


def generate(seed, n, len_lo=256, len_hi=4096, mix=0, edge=0, plain=False):
    """returns (offsets u64[n+1], bytes u8[total], meta[n], ts u64[n])"""
    cfg = GenCfg(seed, n, len_lo, len_hi, mix, edge, 1 if plain else 0)
    total = lib().orc_generate(C.byref(cfg), None, None, None, None)
    offsets = np.zeros(n + 1, dtype=np.uint64)
    data = np.zeros(max(int(total), 1), dtype=np.uint8)
    meta = np.zeros(n, dtype=META_DTYPE)
    ts = np.zeros(n, dtype=np.uint64)
    lib().orc_generate(C.byref(cfg), _p(offsets), _p(data), _p(meta), _p(ts))
    return offsets, data[:int(total)], meta, ts


def pack(contents: list[bytes]):
    offsets = np.zeros(len(contents) + 1, dtype=np.uint64)
    np.cumsum([len(c) for c in contents], out=offsets[1:])
    data = np.frombuffer(b"".join(contents) or b"\0", dtype=np.uint8).copy()
    return offsets, data[:int(offsets[-1])] if offsets[-1] else data[:0]


def process_batch(d: Dict, offsets, data, meta, default_prio=3, nthreads=1):
    n = len(offsets) - 1
    out = np.zeros(n, dtype=RESULT_DTYPE)
    buf = data if len(data) else np.zeros(1, dtype=np.uint8)
    lib().orc_process_batch(d.h, default_prio, _p(offsets), _p(buf), _p(meta), n, _p(out), nthreads)
    return out


def admit_and_order(res, meta, offsets, ts, rules_on=False, max_size=0):
    n = len(res)
    order = np.zeros(max(n, 1), dtype=np.uint32)
    k = lib().orc_admit_and_order(_p(res), _p(meta), _p(offsets), _p(ts), n, 1 if rules_on else 0, max_size, _p(order))
    return order[:k]


def pipeline(d: Dict, offsets, data, meta, ts, default_prio=3, rules_on=False, max_size=0, nthreads=1):
    n = len(offsets) - 1
    out = np.zeros(n, dtype=RESULT_DTYPE)
    order = np.zeros(max(n, 1), dtype=np.uint32)
    buf = data if len(data) else np.zeros(1, dtype=np.uint8)
    k = lib().orc_pipeline(d.h, default_prio, _p(offsets), _p(buf), _p(meta), _p(ts), n,
                           1 if rules_on else 0, max_size, _p(out), _p(order), nthreads)
    return out, order[:k]


def scores(d: Dict, content: bytes):
    sc = (C.c_uint32 * 5)()
    lib().orc_scores(d.h, content, len(content), sc)
    return list(sc)


def analyze(content: bytes):
    w, p, n, q = C.c_uint32(), C.c_int32(), C.c_int32(), C.c_int()
    lib().orc_analyze(content, len(content), C.byref(w), C.byref(p), C.byref(n), C.byref(q))
    return w.value, p.value, n.value, bool(q.value)


def regex_count(pattern: bytes, content: bytes):
    """len(FindAllString(content, -1)) by the C restatement; None when the pattern is outside its subset"""
    h = lib().orc_regex_compile(pattern, len(pattern))
    if not h:
        return None
    try:
        return lib().orc_regex_count(h, content, len(content))
    finally:
        lib().orc_regex_free(h)


def next_backoff_ns(initial_ns, max_ns, factor, retry_count):
    return int(lib().orc_next_backoff_ns(int(initial_ns), int(max_ns), float(factor), int(retry_count)))
