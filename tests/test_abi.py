"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/triage.h
declares, and refuses to compute without a B200 (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re
import subprocess

import pytest

import llm_message_queue_b200 as mq
from tests import hostcheck as hc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "triage.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(triage_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 25
    lib = C.CDLL(mq._capi.build())
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/triage.h but not exported"
    assert set(names) == set(mq._capi._SIGS), "python binding and header disagree"


def test_library_has_sm100a_code_and_tma():
    out = subprocess.run(["cuobjdump", "-lelf", mq._capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    sass = subprocess.run(["cuobjdump", "-sass", mq._capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "UBLKCP" in sass          # cp.async.bulk (1-D TMA) staging of the arena tiles
    assert "UTMALDG" in sass         # cp.async.bulk.tensor.2d: the row-lane scan's window copies
    assert "SYNCS" in sass           # mbarrier


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(mq.TriageError) as ei:
        mq.Engine()
    assert ei.value.status == mq._capi.E_CUDA


def test_product_does_not_reference_the_oracle():
    pkg = os.path.join(ROOT, "llm-message-queue_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cc", ".cuh", ".h", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"oracle[/.]|import oracle|from oracle|liboracle|c_oracle|py_oracle", text), f"{f} uses the oracle"
    out = subprocess.run(["ldd", mq._capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "liboracle" not in out


def test_regex_literal_parser():
    lib = hc.lib()
    def parse(s: bytes):
        ci, n = C.c_int(), C.c_uint32()
        buf = C.create_string_buffer(256)
        rc = lib.hostcheck_parse(s, len(s), C.byref(ci), buf, 256, C.byref(n))
        return rc, bool(ci.value), buf.raw[:n.value]
    assert parse(b"(?i)urgent") == (0, True, b"urgent")
    assert parse(b"right now") == (0, False, b"right now")
    assert parse(rb"a\.b\(c\)") == (0, False, b"a.b(c)")
    assert parse("紧急".encode()) == (0, False, "紧急".encode())
    for bad in (b"urg.nt", b"a|b", b"(urgent)", b"ur+gent", b"[ab]", b"^x", b"x$", rb"\d+", b"", b"(?i)", b"a\\"):
        assert parse(bad)[0] == 4, bad


def test_bordered_literal_detection():
    lib = hc.lib()
    b = lambda s, ci: lib.hostcheck_bordered(s, len(s), ci)  # noqa: E731
    for w in (b"immediate", b"emergency", b"asap", b"right now", b"urgent", b"important", b"priority", b"critical", b"soon"):
        assert b(w, 1) == 0, w   # the nine defaults are border-free
    assert b(b"aa", 0) == 1 and b(b"abab", 0) == 1 and b(b"haha", 0) == 1 and b(b"aA", 1) == 1 and b(b"aA", 0) == 0
