"""CPU-only: the C-ABI library loads and exports every symbol include/mnenc.h declares; host-only entry points work;
GPU entry points fail loudly (no CPU fallback) when no device is present."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from imageexperiments_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "mnenc.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mnenc_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported():
    L = _lib.load()
    names = declared_symbols()
    assert len(names) >= 12
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/mnenc.h but not exported"
    assert sorted(_lib.EXPORTS) == names


def test_no_torch_or_oracle_in_abi_library():
    out = os.popen(f"ldd {_lib.LIB_PATH}").read()
    assert "torch" not in out and "mnoracle" not in out and "mnref" not in out


def test_quant_tables_bad_args():
    L = _lib.load()
    q = np.zeros(96, np.float32)
    assert L.mnenc_quant_tables(64, 8, 3.5, q.ctypes.data_as(C.c_void_p)) < 0
    assert b"K" in L.mnenc_last_error() or L.mnenc_last_error()


def test_create_rejects_bad_arguments():
    L = _lib.load()
    h = C.c_void_p()
    base = np.zeros((510, 64), np.float32)
    rows = np.full(510, 62, np.int32)
    det = np.zeros((62 * 510, 64), np.float32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    assert L.mnenc_create(C.byref(h), 0, 32, 16, p(base), 510, p(rows), p(det), p(det), p(det)) == -1  # block size
    assert L.mnenc_create(C.byref(h), 0, 33, 8, p(base), 510, p(rows), p(det), p(det), p(det)) == -1   # K
    assert L.mnenc_create(C.byref(h), 0, 32, 8, None, 510, p(rows), p(det), p(det), p(det)) == -1        # null


def test_create_auto_reads_the_device_list(monkeypatch):
    """mnenc_create_auto: MNENC_DEVICES must be a comma-separated list of device numbers (or "all"); a malformed list is an
    argument error before any CUDA call, a well-formed one goes on to mnenc_create / mnenc_create_multi."""
    L = _lib.load()
    h = C.c_void_p()
    base = np.zeros((510, 64), np.float32)
    rows = np.full(510, 62, np.int32)
    det = np.zeros((62 * 510, 64), np.float32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for bad in ("0;1", "a", "0,,x", "-1"):
        monkeypatch.setenv("MNENC_DEVICES", bad)
        assert L.mnenc_create_auto(C.byref(h), 32, 8, p(base), 510, p(rows), p(det), p(det), p(det)) == -1, bad
        assert b"MNENC_DEVICES" in L.mnenc_last_error()
    monkeypatch.setenv("MNENC_DEVICES", "0")
    assert L.mnenc_create_auto(C.byref(h), 32, 16, p(base), 510, p(rows), p(det), p(det), p(det)) == -1  # reaches mnenc_create: block size
    assert b"block_size" in L.mnenc_last_error()


def test_gpu_entry_points_fail_loudly_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = _lib.load()
    h = C.c_void_p()
    base = np.zeros((510, 64), np.float32)
    rows = np.full(510, 62, np.int32)
    det = np.zeros((62 * 510, 64), np.float32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = L.mnenc_create(C.byref(h), 0, 32, 8, p(base), 510, p(rows), p(det), p(det), p(det))
    assert rc == -2, "must report MNENC_ERR_CUDA, never fall back to a CPU path"
    assert L.mnenc_last_error()


def test_batch_and_stripe_entry_points_reject_bad_arguments():
    """No context can exist without a GPU, so the batch / stripe entry points must refuse a null context (MNENC_ERR_ARG)
    instead of touching the device; the product has no CPU path behind them."""
    L = _lib.load()
    out, n = C.c_void_p(), C.c_size_t()
    one = (C.c_void_p * 1)()
    sz = (C.c_size_t * 1)(8)
    assert L.mnenc_encode_batch_rgb8(None, 1, one, 0, sz, sz, sz, one, sz) == -1
    rec = np.zeros(99, np.uint32)
    assert L.mnenc_stripe_begin(None, None, 0, 8, 8, 24, 0, 1, 0, 1, rec.ctypes.data_as(C.c_void_p)) == -1
    for f in (L.mnenc_stripe_rle, ):
        assert f(None, rec.ctypes.data_as(C.c_void_p), rec.ctypes.data_as(C.c_void_p)) == -1
    assert L.mnenc_stripe_plan(None, rec.ctypes.data_as(C.c_void_p)) == -1
    assert L.mnenc_stripe_finish(None, C.byref(out), C.byref(n)) == -1
    w, h = C.c_size_t(), C.c_size_t()
    assert L.mnenc_decode_rgb8(None, rec.ctypes.data_as(C.c_void_p), 99, C.byref(out), C.byref(w), C.byref(h)) == -1
    assert L.mnenc_decode_rgb8_device(None, rec.ctypes.data_as(C.c_void_p), 99, rec.ctypes.data_as(C.c_void_p), 99, 3, C.byref(w), C.byref(h)) == -1
    assert L.mnenc_stream_info(rec.ctypes.data_as(C.c_void_p), 99 * 4, None, None, None, None, None) == -5  # not a .mn stream
    L.mnenc_free(None)  # releasing nothing is allowed


def test_python_host_fails_loudly_without_library(monkeypatch, tmp_path):
    """The Python mirror has no fallback either: a missing libmnenc.so is an error that names the build command."""
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "libmnenc.so"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.load()


def test_no_exception_crosses_the_abi():
    """include/mnenc.h: "No exception crosses the ABI" — a host-side std::length_error / bad_alloc comes back as an error code."""
    L = _lib.load()
    buf = (C.c_uint32 * 16)()
    b8 = (C.c_uint8 * 64)()
    u64 = (C.c_uint64 * 2)()
    p = lambda x: C.cast(x, C.c_void_p)  # noqa: E731
    rc = L.mnenc_debug_plan_sequence(p(buf), p(buf), p(buf), 1 << 62, p(buf), p(buf), p(b8), p(buf), p(b8), p(b8), 64, p(u64), p(u64))
    assert rc == -4, rc  # MNENC_ERR_INTERNAL
    text = L.mnenc_last_error().decode()
    assert "exception" in text or "memory" in text, text
