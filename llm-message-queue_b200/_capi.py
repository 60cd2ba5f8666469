"""ctypes binding of libtriage_b200.so (include/triage.h).  The library is the product: if it is
missing or cannot be built, importing this module fails loudly — there is no Python/CPU fallback."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtriage_b200.so")
CSRC = os.path.join(_HERE, "csrc")

OK, E_NOT_FOUND, E_FULL, E_EMPTY, E_PATTERN, E_ARG, E_CUDA = 0, 1, 2, 3, 4, 5, 100

META_DTYPE = np.dtype([("preset", "u1"), ("user_priority", "u1"), ("user_default", "u1"),
                       ("meta_priority", "u1"), ("flags", "u1"), ("reserved", "u1", (11,))])
RESULT_DTYPE = np.dtype([("label", "u1"), ("reason", "u1"), ("sentiment", "i1"), ("question", "u1"),
                         ("accepted", "u1"), ("admit", "u1"), ("flags", "u1"), ("queue_level", "u1"),
                         ("words", "<u4"), ("score_rt", "<u2"), ("score_hi", "<u2")])
RF_ANALYZED, RF_TIE, RF_HAS_ANALYSIS = 1, 2, 4
META_VIP = 1
META_ANALYZE = 2
UP_UNRECOGNISED = 5


class Opts(C.Structure):
    _fields_ = [("device", C.c_int32), ("rules_enabled", C.c_int32), ("max_queue_size", C.c_int64),
                ("max_batch_bytes", C.c_uint64), ("max_batch_msgs", C.c_uint32), ("flags", C.c_uint32)]


class QStats(C.Structure):
    _fields_ = [("pending", C.c_int64), ("processing", C.c_int64), ("completed", C.c_int64), ("failed", C.c_int64)]


class Timings(C.Structure):
    _fields_ = [("scan_ms", C.c_float), ("finalize_ms", C.c_float), ("order_ms", C.c_float), ("total_ms", C.c_float),
                ("h2d_ms", C.c_float), ("d2h_ms", C.c_float), ("launches", C.c_uint32), ("order_passes", C.c_uint32),
                ("scan_bytes", C.c_uint64), ("scan_variant", C.c_uint32), ("scan_row_bytes", C.c_uint32), ("scan_warps", C.c_uint32),
                ("reserved", C.c_uint32)]

SCAN_VARIANTS = {0: "none", 1: "scan_kernel (packed arena, CTA-wide tiles)", 2: "scan_padded_kernel (padded arena, mini-tiles)",
                 3: "scan_rows_kernel (padded arena, row lanes, 2-D TMA windows of 80 B x 32 rows)",
                 4: "scan_padded_kernel<wide> (class map in shared memory, transitions in L2)",
                 5: "scan_pair_kernel (padded arena, row lanes fed by 2-D TMA, filter automaton over class pairs: 8 lookups per 16 bytes, "
                    "SIMD word count, flagged groups walked exactly)",
                 6: "scan_stream_wide_kernel (wide table in L2/L1, rows streamed from global memory, several rows per lane walked interleaved)"}


class TriageError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"triage status {status}: {message}")
        self.status = status
        self.message = message


def build(force: bool = False) -> str:
    """nvcc -gencode arch=compute_100a,code=sm_100a build of the in-tree library (cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cc", ".cuh", ".h"))]
    srcs.append(os.path.join(os.path.dirname(_HERE), "include", "triage.h"))
    stale = not os.path.exists(LIB_PATH) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", CSRC, "-s"] + (["-B"] if force else []), stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None

# every symbol include/triage.h declares (tests check the export list against the header)
_SIGS = {
    "triage_create": (C.c_int32, [C.POINTER(Opts), C.POINTER(C.c_void_p)]),
    "triage_destroy": (None, [C.c_void_p]),
    "triage_last_error": (C.c_char_p, []),
    "triage_dict_reset": (C.c_int32, [C.c_void_p, C.c_int32]),
    "triage_dict_add_pattern": (C.c_int32, [C.c_void_p, C.c_int32, C.c_char_p, C.c_uint32]),
    "triage_dict_add_literal": (C.c_int32, [C.c_void_p, C.c_int32, C.c_char_p, C.c_uint32, C.c_int32]),
    "triage_dict_commit": (C.c_int32, [C.c_void_p]),
    "triage_dict_count": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(C.c_uint32)]),
    "triage_set_default_priority": (C.c_int32, [C.c_void_p, C.c_int32]),
    "triage_dict_info": (C.c_int32, [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "triage_host_alloc": (C.c_int32, [C.POINTER(C.c_void_p), C.c_uint64]),
    "triage_host_free": (C.c_int32, [C.c_void_p]),
    "triage_submit": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32,
                                  C.c_void_p, C.POINTER(C.c_uint64)]),
    "triage_process": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
    "triage_submit_device": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                         C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32)]),
    "triage_pad_byte": (C.c_int32, [C.c_void_p, C.POINTER(C.c_int32)]),
    "triage_pack_padded": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.POINTER(C.c_uint64)]),
    "triage_submit_padded": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_uint32, C.c_void_p, C.POINTER(C.c_uint64)]),
    "triage_process_padded": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
    "triage_prefetch": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]),
    "triage_submit_device_padded": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                                                C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32)]),
    "triage_submit_device_static": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                                C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32)]),
    "triage_queue_create": (C.c_int32, [C.c_void_p, C.c_int64, C.POINTER(C.c_int32)]),
    "triage_queue_delete": (C.c_int32, [C.c_void_p, C.c_int32]),
    "triage_set_limits": (C.c_int32, [C.c_void_p, C.c_int64, C.c_int32]),
    "triage_pop": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(C.c_uint64)]),
    "triage_peek": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(C.c_uint64)]),
    "triage_drain": (C.c_int32, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32)]),
    "triage_push": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.c_uint64, C.POINTER(C.c_uint64)]),
    "triage_queue_size": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(C.c_uint64)]),
    "triage_stats": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(QStats)]),
    "triage_complete": (C.c_int32, [C.c_void_p, C.c_int32]),
    "triage_fail": (C.c_int32, [C.c_void_p, C.c_int32]),
    "triage_set_user_priority": (C.c_int32, [C.c_void_p, C.c_char_p, C.c_uint32, C.c_int32]),
    "triage_ingest_json": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]),
    "triage_emit_analysis_json": (C.c_int32, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p, C.POINTER(C.c_uint64)]),
    "triage_drain_queue": (C.c_int32, [C.c_void_p, C.c_int32, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint32)]),
    "triage_complete_batch": (C.c_int32, [C.c_void_p, C.c_int32, C.c_uint32]),
    "triage_fail_batch": (C.c_int32, [C.c_void_p, C.c_int32, C.c_uint32]),
    "triage_requeue": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint32)]),
    "triage_schedule": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]),
    "triage_release_due": (C.c_int32, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "triage_delayed_peek": (C.c_int32, [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "triage_delayed_size": (C.c_int32, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "triage_backoff_ns": (C.c_int64, [C.c_int64, C.c_int64, C.c_double, C.c_int32]),
    "triage_last_timings": (C.c_int32, [C.c_void_p, C.POINTER(Timings)]),
    "triage_event_record": (C.c_int32, [C.c_void_p, C.c_int32]),
    "triage_event_elapsed_ms": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_float)]),
    "triage_comm_unique_id": (C.c_int32, [C.c_void_p]),
    "triage_comm_init": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "triage_merge_allgather": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "triage_merge_allgather_async": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]),
    "triage_merge_wait": (C.c_int32, [C.c_void_p, C.POINTER(C.c_uint64)]),
}


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)   # AttributeError if the library does not export it
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _check(rc: int):
    if rc != OK:
        raise TriageError(rc, (lib().triage_last_error() or b"").decode(errors="replace"))


def _p(a):
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p)


def pack_contents(contents):
    """list[bytes] -> (offsets u64[n+1], bytes u8[total])"""
    offsets = np.zeros(len(contents) + 1, dtype=np.uint64)
    if contents:
        np.cumsum([len(c) for c in contents], out=offsets[1:])
    data = np.frombuffer(b"".join(contents), dtype=np.uint8).copy() if offsets[-1] else np.zeros(0, np.uint8)
    return offsets, data


def pack_padded(offsets, data, pad: int, out_bytes=None):
    n = len(offsets) - 1
    dst_off = np.zeros(n + 1, np.uint64)
    dst_len = np.zeros(max(n, 1), np.uint32)
    total = C.c_uint64()
    src = data if len(data) else np.zeros(1, np.uint8)
    _check(lib().triage_pack_padded(_p(offsets), _p(src), n, pad, _p(dst_off), _p(dst_len), None, C.byref(total)))
    if out_bytes is None:
        out_bytes = np.empty(max(total.value, 16), np.uint8)
    elif len(out_bytes) < total.value:
        raise ValueError("out_bytes too small")
    _check(lib().triage_pack_padded(_p(offsets), _p(src), n, pad, _p(dst_off), _p(dst_len), _p(out_bytes), C.byref(total)))
    return dst_off, dst_len[:n], out_bytes[:total.value]


class Engine:
    """One triage_engine (one GPU, one set of level queues)."""

    def __init__(self, device: int = 0, rules_enabled: bool = False, max_queue_size: int = 0):
        self._h = C.c_void_p()
        opts = Opts(device, 1 if rules_enabled else 0, max_queue_size, 0, 0, 0)
        _check(lib().triage_create(C.byref(opts), C.byref(self._h)))

    def close(self):
        if self._h:
            lib().triage_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- dictionary -------------------------------------------------------
    def dict_reset(self, with_defaults=True):
        _check(lib().triage_dict_reset(self._h, 1 if with_defaults else 0))

    def add_pattern(self, prio: int, regex: str | bytes):
        b = regex.encode("utf-8") if isinstance(regex, str) else regex
        _check(lib().triage_dict_add_pattern(self._h, int(prio), b, len(b)))

    def add_literal(self, prio: int, literal: bytes, ci: bool):
        _check(lib().triage_dict_add_literal(self._h, int(prio), literal, len(literal), 1 if ci else 0))

    def commit(self):
        _check(lib().triage_dict_commit(self._h))

    def dict_count(self, prio: int) -> int:
        n = C.c_uint32()
        _check(lib().triage_dict_count(self._h, int(prio), C.byref(n)))
        return n.value

    def dict_info(self):
        a, b, c = C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(lib().triage_dict_info(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return dict(states=a.value, warm_bytes=b.value, table_bytes=c.value)

    def set_default_priority(self, prio: int):
        _check(lib().triage_set_default_priority(self._h, int(prio)))

    def set_limits(self, max_queue_size: int, rules_enabled: bool):
        _check(lib().triage_set_limits(self._h, int(max_queue_size), 1 if rules_enabled else 0))

    # -- hot path -----------------------------------------------------------
    def process(self, offsets, data, meta):
        """ProcessMessage for a batch given as a packed arena; dictionaries too large for the dense table are served
        through the padded layout (re-packed here)."""
        n = len(offsets) - 1
        out = np.zeros(n, dtype=RESULT_DTYPE)
        if n and self.dict_info()["states"] > 256:
            poff, plen, pbytes = self.pack_padded(offsets, data)
            _check(lib().triage_process_padded(self._h, _p(poff), _p(plen), _p(pbytes), _p(meta), n, _p(out)))
            return out
        buf = data if len(data) else np.zeros(16, np.uint8)
        _check(lib().triage_process(self._h, _p(offsets), _p(buf), _p(meta), n, _p(out)))
        return out

    def submit(self, offsets, data, meta, ts=None, target_queue=0):
        n = len(offsets) - 1
        out = np.zeros(n, dtype=RESULT_DTYPE)
        base = C.c_uint64()
        buf = data if len(data) else np.zeros(16, np.uint8)
        _check(lib().triage_submit(self._h, int(target_queue), _p(offsets), _p(buf), _p(meta), _p(ts), n, _p(out), C.byref(base)))
        return out, base.value

    def pad_byte(self) -> int:
        v = C.c_int32()
        _check(lib().triage_pad_byte(self._h, C.byref(v)))
        return v.value

    def pack_padded(self, offsets, data, out_bytes=None):
        """packed (offsets, bytes) -> padded layout (offsets u64[n+1], lengths u32[n], bytes u8[total]).  out_bytes: an
        optional preallocated (e.g. pinned) uint8 array to fill."""
        return pack_padded(offsets, data, self.pad_byte(), out_bytes)

    def submit_padded(self, offsets, lengths, data, meta, ts=None, target_queue=0):
        n = len(lengths)
        out = np.zeros(n, dtype=RESULT_DTYPE)
        base = C.c_uint64()
        _check(lib().triage_submit_padded(self._h, int(target_queue), _p(offsets), _p(lengths), _p(data), _p(meta), _p(ts), n, _p(out),
                                          C.byref(base)))
        return out, base.value

    def prefetch(self, offsets, lengths, data, meta, ts=None):
        """start the host->device copy of a batch; a later submit / submit_padded with the same arrays skips the upload.
        lengths=None: packed arena.  The arrays must stay alive and unchanged until that submit returns."""
        _check(lib().triage_prefetch(self._h, _p(offsets), _p(lengths), _p(data), _p(meta), _p(ts), len(meta)))

    def submit_device_padded(self, offsets_ptr, lengths_ptr, bytes_ptr, total, meta_ptr, ts_ptr, n, results_ptr, order_ptr) -> int:
        adm = C.c_uint32()
        _check(lib().triage_submit_device_padded(self._h, offsets_ptr, lengths_ptr, bytes_ptr, int(total), meta_ptr, ts_ptr, int(n),
                                                 results_ptr, order_ptr, C.byref(adm)))
        return adm.value

    def submit_device(self, offsets_ptr, bytes_ptr, total, meta_ptr, ts_ptr, n, results_ptr, order_ptr) -> int:
        adm = C.c_uint32()
        _check(lib().triage_submit_device(self._h, offsets_ptr, bytes_ptr, int(total), meta_ptr, ts_ptr, int(n), results_ptr,
                                          order_ptr, C.byref(adm)))
        return adm.value

    def submit_device_static(self, offsets_ptr, bytes_ptr, total, meta_ptr, ts_ptr, n, results_ptr, order_ptr) -> int:
        """triage_submit_device without any device allocation: scratch lives in a fixed 3 MiB array inside the library
        (<= 65536 messages, <= 256 MiB arena, dictionary committed beforehand)."""
        adm = C.c_uint32()
        _check(lib().triage_submit_device_static(self._h, offsets_ptr, bytes_ptr, int(total), meta_ptr, ts_ptr, int(n), results_ptr,
                                                 order_ptr, C.byref(adm)))
        return adm.value

    def timings(self) -> dict:
        t = Timings()
        _check(lib().triage_last_timings(self._h, C.byref(t)))
        return {k: getattr(t, k) for k, _ in Timings._fields_}

    def event_record(self, slot: int):
        _check(lib().triage_event_record(self._h, slot))

    def event_elapsed_ms(self, a: int, b: int) -> float:
        ms = C.c_float()
        _check(lib().triage_event_elapsed_ms(self._h, a, b, C.byref(ms)))
        return ms.value

    def host_array(self, nbytes: int, dtype=np.uint8):
        """numpy array over pinned host memory (triage_host_alloc); freed with the engine's process."""
        p = C.c_void_p()
        _check(lib().triage_host_alloc(C.byref(p), int(nbytes)))
        buf = (C.c_uint8 * int(nbytes)).from_address(p.value)
        return np.frombuffer(buf, dtype=np.uint8).view(dtype)

    # -- queues -------------------------------------------------------------
    def queue_create(self, max_size=0) -> int:
        q = C.c_int32()
        _check(lib().triage_queue_create(self._h, int(max_size), C.byref(q)))
        return q.value

    def queue_delete(self, q: int):
        _check(lib().triage_queue_delete(self._h, int(q)))

    def push(self, q: int, item_priority: int, enqueue_ns: int = 0) -> int:
        h = C.c_uint64()
        _check(lib().triage_push(self._h, int(q), int(item_priority), int(enqueue_ns), C.byref(h)))
        return h.value

    def pop(self, q: int) -> int:
        h = C.c_uint64()
        _check(lib().triage_pop(self._h, int(q), C.byref(h)))
        return h.value

    def peek(self, q: int) -> int:
        h = C.c_uint64()
        _check(lib().triage_peek(self._h, int(q), C.byref(h)))
        return h.value

    def drain(self, max_items: int):
        handles = np.zeros(max(max_items, 1), np.uint64)
        levels = np.zeros(max(max_items, 1), np.uint8)
        k = C.c_uint32()
        _check(lib().triage_drain(self._h, int(max_items), _p(handles), _p(levels), C.byref(k)))
        return handles[:k.value], levels[:k.value]

    def queue_size(self, q: int) -> int:
        s = C.c_uint64()
        _check(lib().triage_queue_size(self._h, int(q), C.byref(s)))
        return s.value

    def stats(self, q: int) -> dict:
        st = QStats()
        _check(lib().triage_stats(self._h, int(q), C.byref(st)))
        return dict(pending=st.pending, processing=st.processing, completed=st.completed, failed=st.failed)

    def complete(self, q: int):
        _check(lib().triage_complete(self._h, int(q)))

    def fail(self, q: int):
        _check(lib().triage_fail(self._h, int(q)))

    # -- wire format: JSON in, analysis JSON out ------------------------------
    def set_user_priority(self, user_id: str | bytes, prio: int):
        b = user_id.encode("utf-8") if isinstance(user_id, str) else user_id
        _check(lib().triage_set_user_priority(self._h, b, len(b), int(prio)))

    def ingest_json(self, bodies: list[bytes]):
        """request bodies -> (offsets, lengths, arena bytes, meta, status) in the padded layout, ready for submit_padded"""
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
        _check(lib().triage_ingest_json(self._h, _p(blob), _p(boff), n, None, 0, _p(off), _p(lens), _p(meta), _p(status), C.byref(total)))
        arena = np.empty(max(total.value, 16), np.uint8)
        _check(lib().triage_ingest_json(self._h, _p(blob), _p(boff), n, _p(arena), len(arena), _p(off), _p(lens), _p(meta), _p(status), C.byref(total)))
        return off, lens[:n], arena[:total.value], meta[:n], status[:n]

    @staticmethod
    def emit_analysis_json(results) -> list[bytes]:
        n = len(results)
        res = np.ascontiguousarray(results)
        need = C.c_uint64()
        _check(lib().triage_emit_analysis_json(_p(res), n, None, 0, None, C.byref(need)))
        out = np.zeros(max(need.value, 1), np.uint8)
        offs = np.zeros(n + 1, np.uint64)
        _check(lib().triage_emit_analysis_json(_p(res), n, _p(out), len(out), _p(offs), C.byref(need)))
        raw = out.tobytes()
        return [raw[int(offs[i]):int(offs[i + 1])] for i in range(n)]

    # -- batch consumer, retries, delayed items --------------------------------
    def drain_queue(self, q: int, max_items: int):
        """BatchPopMessages: handles popped from one queue, until it is empty or max_items"""
        handles = np.zeros(max(max_items, 1), np.uint64)
        k = C.c_uint32()
        _check(lib().triage_drain_queue(self._h, int(q), int(max_items), _p(handles), C.byref(k)))
        return handles[:k.value]

    def complete_batch(self, q: int, k: int):
        _check(lib().triage_complete_batch(self._h, int(q), int(k)))

    def fail_batch(self, q: int, k: int):
        _check(lib().triage_fail_batch(self._h, int(q), int(k)))

    def requeue(self, q: int, handles, item_priority, enqueue_ns=None):
        """PushMessage for messages that already have handles (retries, dead-letter requeue) -> accepted u8[n]"""
        handles = np.ascontiguousarray(handles, np.uint64)
        prio = np.ascontiguousarray(item_priority, np.int32)
        ts = None if enqueue_ns is None else np.ascontiguousarray(enqueue_ns, np.uint64)
        acc = np.zeros(max(len(handles), 1), np.uint8)
        k = C.c_uint32()
        _check(lib().triage_requeue(self._h, int(q), _p(handles), _p(prio), _p(ts), len(handles), _p(acc), C.byref(k)))
        return acc[:len(handles)]

    def schedule(self, q: int, handles, item_priority, ready_at_ns):
        handles = np.ascontiguousarray(handles, np.uint64)
        prio = np.ascontiguousarray(item_priority, np.int32)
        ready = np.ascontiguousarray(ready_at_ns, np.uint64)
        _check(lib().triage_schedule(self._h, int(q), _p(handles), _p(prio), _p(ready), len(handles)))

    def release_due(self, now_ns: int = 0, max_items: int = 0):
        """-> (handles released in ReadyAt order, their queues, number a full or deleted queue turned away)"""
        cap = max_items if max_items else max(self.delayed_size(), 1)
        handles = np.zeros(cap, np.uint64)
        queues = np.zeros(cap, np.int32)
        k, r = C.c_uint32(), C.c_uint32()
        _check(lib().triage_release_due(self._h, int(now_ns), _p(handles), _p(queues), int(max_items), C.byref(k), C.byref(r)))
        return handles[:k.value], queues[:k.value], r.value

    def delayed_peek(self):
        h, t = C.c_uint64(), C.c_uint64()
        _check(lib().triage_delayed_peek(self._h, C.byref(h), C.byref(t)))
        return h.value, t.value

    def delayed_size(self) -> int:
        n = C.c_uint64()
        _check(lib().triage_delayed_size(self._h, C.byref(n)))
        return n.value

    # -- multi-GPU ------------------------------------------------------------
    @staticmethod
    def comm_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        _check(lib().triage_comm_unique_id(buf))
        return buf.raw

    def comm_init(self, nranks: int, rank: int, uid: bytes):
        _check(lib().triage_comm_init(self._h, nranks, rank, C.create_string_buffer(uid, 128)))

    def merge_allgather_async(self, rank_base: np.ndarray, order_global_ptr: int, capacity: int) -> None:
        """start the one-exchange merge on the merge stream (engine-stamped enqueue times only); merge_wait() completes it"""
        self._rank_base_keep = np.ascontiguousarray(rank_base, np.uint64)
        _check(lib().triage_merge_allgather_async(self._h, _p(self._rank_base_keep), order_global_ptr, int(capacity)))

    def merge_wait(self) -> int:
        tot = C.c_uint64()
        _check(lib().triage_merge_wait(self._h, C.byref(tot)))
        return tot.value

    def merge_allgather(self, rank_base: np.ndarray, order_global_ptr: int, capacity: int) -> int:
        tot = C.c_uint64()
        _check(lib().triage_merge_allgather(self._h, _p(rank_base), order_global_ptr, int(capacity), C.byref(tot)))
        return tot.value


def backoff_ns(initial_ns: int, max_ns: int, factor: float, retry_count: int) -> int:
    """ExponentialBackoff.NextBackoff (worker.go:274-290) in nanoseconds"""
    return int(lib().triage_backoff_ns(int(initial_ns), int(max_ns), float(factor), int(retry_count)))
