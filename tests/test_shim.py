"""The C++ drop-in shim (shim/encode_image_fast_shim.cpp): reference signature of compressed::encodeImageFast, dictionary
recovered through the opaque Dynamic callbacks, GPU encode through the C ABI. Output must equal the reference's own
encodeImageFast byte for byte."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "oracle", "_ref", "libmnshim.so")


def test_shim_library_links_against_the_c_abi():
    if not os.path.exists(SHIM):
        pytest.skip("shim not built (needs the reference headers; __graft_entry__.build())")
    out = os.popen(f"ldd {SHIM}").read()
    assert "libmnenc.so" in out


@pytest.mark.gpu
def test_shim_equals_reference_encoder(reflib):
    if not os.path.exists(SHIM):
        pytest.skip("shim not built")
    from imageexperiments_b200.synth import synthetic_rgb
    L = C.CDLL(SHIM)
    L.mnenc_shim_selftest.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_double, C.c_void_p, C.c_void_p]
    img = synthetic_rgb(136, 96, 31)
    out, n = C.c_void_p(), C.c_size_t()
    assert L.mnenc_shim_selftest(img.ctypes.data_as(C.c_void_p), 136, 96, 3.5, C.byref(out), C.byref(n)) == 0
    got = C.string_at(out, n.value)
    ctx = reflib.create(32, 8, 3.5)
    assert got == reflib.encode(ctx, img)
    reflib.destroy(ctx)


@pytest.mark.gpu
def test_shim_equals_reference_decoder(reflib):
    """mnenc_shim::decodeImageFast (reference signature) -> same pixels as compressed::decodeImageFast."""
    if not os.path.exists(SHIM):
        pytest.skip("shim not built")
    from imageexperiments_b200.synth import synthetic_rgb
    L = C.CDLL(SHIM)
    L.mnenc_shim_selftest_decode.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    ctx = reflib.create(32, 8, 3.5)
    data = reflib.encode(ctx, synthetic_rgb(136, 96, 31))
    reflib.destroy(ctx)
    out, w, h = C.c_void_p(), C.c_size_t(), C.c_size_t()
    assert L.mnenc_shim_selftest_decode(data, len(data), C.byref(out), C.byref(w), C.byref(h)) == 0
    want = reflib.decode(data)
    assert (h.value, w.value) == want.shape[:2]
    got = np.frombuffer(C.string_at(out, 3 * w.value * h.value), np.uint8).reshape(want.shape)
    assert (got == want).all()
    assert L.mnenc_shim_selftest_decode(data[:40], 40, C.byref(out), C.byref(w), C.byref(h)) == -1  # thrown range_error, as the reference
