"""Test helper: builds and binds tests/csrc/libscan_host_check.so (g++ build of the product's
dictionary compiler + per-lane scan body).  TEST HARNESS — not a product code path."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = [os.path.join(HERE, "csrc", "scan_host_check.cc"),
       os.path.join(ROOT, "llm-message-queue_b200", "csrc", "dfa_build.cc"),
       os.path.join(ROOT, "llm-message-queue_b200", "csrc", "regex_vm.cc"),
       os.path.join(ROOT, "llm-message-queue_b200", "csrc", "ingest.cc")]
DEPS = SRC + [os.path.join(ROOT, "llm-message-queue_b200", "csrc", f)
              for f in ("dfa_build.h", "scan_core.cuh", "scan_wide.cuh", "scan_pair.cuh", "regex_vm.h", "regex_walk.cuh", "ingest.h")]
LIB = os.path.join(HERE, "csrc", "libscan_host_check.so")

DEFAULT_DICT = [(1, True, b"immediate"), (1, True, b"emergency"), (1, True, b"asap"), (1, True, b"right now"),
                (2, True, b"urgent"), (2, True, b"important"), (2, True, b"priority"), (2, True, b"critical"),
                (2, True, b"soon")]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in DEPS):
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", LIB] + SRC)
        _lib = C.CDLL(LIB)
        _lib.hostcheck_scan.restype = C.c_int
    return _lib


def scan(dict_entries, offsets, data, nt=64, chunk=48, use_abs_list=False, force_slow=False):
    """returns (counters[n,8], info dict) or raises ValueError(build error)"""
    n = len(offsets) - 1
    prios = np.array([d[0] for d in dict_entries], np.uint8)
    cis = np.array([1 if d[1] else 0 for d in dict_entries], np.uint8)
    lens = np.array([len(d[2]) for d in dict_entries], np.uint32)
    lits = np.frombuffer(b"".join(d[2] for d in dict_entries) or b"\0", np.uint8)
    counters = np.zeros((max(n, 1), 8), np.uint32)
    info = np.zeros(6, np.uint32)
    err = C.create_string_buffer(256)
    buf = data if len(data) else np.zeros(1, np.uint8)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = lib().hostcheck_scan(p(prios), p(cis), p(lens), p(lits), len(dict_entries), p(offsets), p(buf), n,
                              nt, chunk, (1 if use_abs_list else 0) | (2 if force_slow else 0), p(counters), p(info), err, 256)
    if rc != 0:
        raise ValueError(err.value.decode())
    counters[:, 7] &= 1   # bit 1 = overlap-risk flag (self-overlapping keywords), not part of the analysis
    return counters[:n], dict(states=int(info[0]), unminimised=int(info[1]), warm=int(info[2]), events=int(info[3]),
                                head_pad=int(np.int32(info[4])), tail_pad=int(np.int32(info[5])))


def scan_padded(dict_entries, offsets, data, chunk=80, use_abs_list=False):
    """same counters through the padded-arena lane (engine's preferred layout)"""
    n = len(offsets) - 1
    prios = np.array([d[0] for d in dict_entries], np.uint8)
    cis = np.array([1 if d[1] else 0 for d in dict_entries], np.uint8)
    lens = np.array([len(d[2]) for d in dict_entries], np.uint32)
    lits = np.frombuffer(b"".join(d[2] for d in dict_entries) or b"\0", np.uint8)
    counters = np.zeros((max(n, 1), 8), np.uint32)
    info = np.zeros(6, np.uint32)
    err = C.create_string_buffer(256)
    buf = data if len(data) else np.zeros(1, np.uint8)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = lib().hostcheck_scan_padded(p(prios), p(cis), p(lens), p(lits), len(dict_entries), p(offsets), p(buf), n,
                                     chunk, 1 if use_abs_list else 0, p(counters), p(info), err, 256)
    if rc != 0:
        raise ValueError(err.value.decode())
    risk = (counters[:n, 7] & 2) != 0
    counters[:, 7] &= 1
    return counters[:n], dict(states=int(info[0]), reset_pad=int(np.int32(info[1])), risk=risk)


def scan_wide(dict_entries, offsets, data, chunk=80):
    """counters through the wide-table lane (large dictionaries) + exact recount of flagged messages"""
    n = len(offsets) - 1
    prios = np.array([d[0] for d in dict_entries], np.uint8)
    cis = np.array([1 if d[1] else 0 for d in dict_entries], np.uint8)
    lens = np.array([len(d[2]) for d in dict_entries], np.uint32)
    lits = np.frombuffer(b"".join(d[2] for d in dict_entries) or b"\0", np.uint8)
    counters = np.zeros((max(n, 1), 8), np.uint32)
    info = np.zeros(6, np.uint32)
    err = C.create_string_buffer(256)
    buf = data if len(data) else np.zeros(1, np.uint8)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = lib().hostcheck_scan_wide(p(prios), p(cis), p(lens), p(lits), len(dict_entries), p(offsets), p(buf), n, chunk,
                                   p(counters), p(info), err, 256)
    if rc != 0:
        raise ValueError(err.value.decode())
    return counters[:n], dict(states=int(info[0]), reset_pad=int(np.int32(info[1])), wide=bool(info[2]), classes=int(info[3]),
                              warm=int(info[4]), events=int(info[5]))


def scan_long(dict_entries, offsets, data, chunk_long=320, window=80):
    """counters through the long-lane walk (chunk_long bytes per lane, streamed in windows)"""
    n = len(offsets) - 1
    prios = np.array([d[0] for d in dict_entries], np.uint8)
    cis = np.array([1 if d[1] else 0 for d in dict_entries], np.uint8)
    lens = np.array([len(d[2]) for d in dict_entries], np.uint32)
    lits = np.frombuffer(b"".join(d[2] for d in dict_entries) or b"\0", np.uint8)
    counters = np.zeros((max(n, 1), 8), np.uint32)
    err = C.create_string_buffer(256)
    buf = data if len(data) else np.zeros(1, np.uint8)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = lib().hostcheck_scan_long(p(prios), p(cis), p(lens), p(lits), len(dict_entries), p(offsets), p(buf), n, chunk_long, window,
                                   p(counters), err, 256)
    if rc != 0:
        raise ValueError(err.value.decode())
    return counters[:n]


def scan_stream(dict_entries, offsets, data, chunk_long=512, nch=2, wide=False):
    """counters through the streaming walk (nch interleaved chains of chunk_long bytes per lane); wide: the wide table
    format with events applied where they are met (scan_stream_wide_kernel) + the exact recount of flagged messages"""
    n = len(offsets) - 1
    prios = np.array([d[0] for d in dict_entries], np.uint8)
    cis = np.array([1 if d[1] else 0 for d in dict_entries], np.uint8)
    lens = np.array([len(d[2]) for d in dict_entries], np.uint32)
    lits = np.frombuffer(b"".join(d[2] for d in dict_entries) or b"\0", np.uint8)
    counters = np.zeros((max(n, 1), 8), np.uint32)
    err = C.create_string_buffer(256)
    buf = data if len(data) else np.zeros(1, np.uint8)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    fn = lib().hostcheck_scan_stream_wide if wide else lib().hostcheck_scan_stream
    rc = fn(p(prios), p(cis), p(lens), p(lits), len(dict_entries), p(offsets), p(buf), n, chunk_long, nch, p(counters), err, 256)
    if rc != 0:
        raise ValueError(err.value.decode())
    return counters[:n]


def regex_count(pattern: bytes, content: bytes):
    """len(FindAllString) of the engine's regex subset through the host build of the device walk; ValueError(reason) when
    the pattern is outside the subset"""
    cnt = C.c_uint32()
    err = C.create_string_buffer(256)
    rc = lib().hostcheck_regex_count(pattern, len(pattern), content if content else b"\0", len(content), C.byref(cnt), err, 256)
    if rc != 0:
        raise ValueError(err.value.decode())
    return cnt.value


def scan_pair(dict_entries, offsets, data, chunk_long=320):
    """counters through the two-bytes-per-lookup lane (filter automaton + SIMD word count + exact walk of flagged groups)
    -> (counters[n,8], info dict); ValueError when the dictionary has no pair filter"""
    n = len(offsets) - 1
    prios = np.array([d[0] for d in dict_entries], np.uint8)
    cis = np.array([1 if d[1] else 0 for d in dict_entries], np.uint8)
    lens = np.array([len(d[2]) for d in dict_entries], np.uint32)
    lits = np.frombuffer(b"".join(d[2] for d in dict_entries) or b"\0", np.uint8)
    counters = np.zeros((max(n, 1), 8), np.uint32)
    info = np.zeros(5, np.uint64)
    err = C.create_string_buffer(256)
    buf = data if len(data) else np.zeros(1, np.uint8)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = lib().hostcheck_scan_pair(p(prios), p(cis), p(lens), p(lits), len(dict_entries), p(offsets), p(buf), n, chunk_long,
                                   p(counters), p(info), err, 256)
    if rc != 0:
        raise ValueError(err.value.decode())
    return counters[:n], dict(filter_states=int(info[0]) & 0xFFFFFFFF, exact_classes=int(info[0]) >> 32, mul=int(info[1]), shift=int(info[2]), flagged=int(info[3]), groups=int(info[4]))


META_DTYPE = np.dtype([("preset", "u1"), ("user_priority", "u1"), ("user_default", "u1"), ("meta_priority", "u1"), ("flags", "u1"),
                       ("reserved", "u1", (11,))])
RESULT_DTYPE = np.dtype([("label", "u1"), ("reason", "u1"), ("sentiment", "i1"), ("question", "u1"), ("accepted", "u1"), ("admit", "u1"),
                         ("flags", "u1"), ("queue_level", "u1"), ("words", "<u4"), ("score_rt", "<u2"), ("score_hi", "<u2")])


def ingest(bodies, pad=9, user=None, user_prio=0):
    """host build of triage_ingest_json's decoder -> (contents list[bytes], meta, status)"""
    n = len(bodies)
    boff = np.zeros(n + 1, np.uint64)
    if n:
        np.cumsum([len(b) for b in bodies], out=boff[1:])
    blob = np.frombuffer(b"".join(bodies) or b"\0", np.uint8)
    off = np.zeros(n + 1, np.uint64)
    lens = np.zeros(max(n, 1), np.uint32)
    meta = np.zeros(max(n, 1), META_DTYPE)
    status = np.zeros(max(n, 1), np.uint8)
    total = C.c_uint64()
    err = C.create_string_buffer(256)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    u = user.encode() if user else None
    rc = lib().hostcheck_ingest(p(blob), p(boff), n, pad, u, user_prio, None, 0, p(off), p(lens), p(meta), p(status), C.byref(total), err, 256)
    if rc:
        raise ValueError(err.value.decode())
    arena = np.zeros(max(total.value, 16), np.uint8)
    rc = lib().hostcheck_ingest(p(blob), p(boff), n, pad, u, user_prio, p(arena), len(arena), p(off), p(lens), p(meta), p(status), C.byref(total), err, 256)
    if rc:
        raise ValueError(err.value.decode())
    raw = arena.tobytes()
    contents = [raw[int(off[i]):int(off[i]) + int(lens[i])] for i in range(n)]
    for i in range(n):      # the layout contract of the padded arena
        assert off[i] % 16 == 0 and 1 <= int(off[i + 1]) - int(off[i]) - int(lens[i]) <= 16
        assert set(raw[int(off[i]) + int(lens[i]):int(off[i + 1])]) == {pad}
    return contents, meta[:n], status[:n]


def emit(results):
    res = np.ascontiguousarray(results)
    n = len(res)
    lib().hostcheck_emit.restype = C.c_uint64
    need = lib().hostcheck_emit(res.ctypes.data_as(C.c_void_p), n, None, 0, None)
    out = np.zeros(max(int(need), 1), np.uint8)
    offs = np.zeros(n + 1, np.uint64)
    lib().hostcheck_emit(res.ctypes.data_as(C.c_void_p), n, out.ctypes.data_as(C.c_void_p), len(out), offs.ctypes.data_as(C.c_void_p))
    raw = out.tobytes()
    return [raw[int(offs[i]):int(offs[i + 1])] for i in range(n)]
