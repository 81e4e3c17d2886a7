"""GPU test of mnenc_scatter_streams_static, the one entry point for callers that are not allowed to allocate GPU memory
(a plugin inside a program that owns all of HBM, code on a path where an allocator call would stall the device, a process
whose memory pool took the whole GPU at start-up). It keeps its working data in a fixed array inside the library instead of
a cudaMalloc'ed context buffer, so the things to check are: (1) it produces exactly what the ordinary, allocating path
produces — the 6K symbol streams that mnenc_debug_parse reads back out of the file mnenc_encode_rgb8 wrote (integer data:
the comparison is exact, like every other byte/symbol test here); (2) leftovers of an earlier call in that shared array
do not leak into the next one — the calls below run back to back with different image sizes, large after small after large;
(3) it really allocates nothing once set up; (4) it refuses what it cannot do (too many blocks, a host pointer, an output
buffer that is too small) with an error text instead of a fault.

This file sorts after every other test file on purpose (pytest -x stops at the first failure). It needs nothing from
the tests before it and passes on its own in a fresh process as well."""
import ctypes as C

import numpy as np
import pytest

import imageexperiments_b200 as mn
from imageexperiments_b200 import _lib
from imageexperiments_b200.synth import synthetic_rgb

pytestmark = pytest.mark.gpu

K = 32
MAX_BLOCKS = 1 << 20  # documented limit of the entry point (include/mnenc.h)


@pytest.fixture(scope="module")
def ctx(dictionary):
    c = mn.createCompressionContextFast(K, 8, 3.5, dictionary=dictionary, device=0)
    yield c
    c.close()


def reference_streams(ctx, img):
    """The ordinary path: encode through mnenc_encode_rgb8, then the symbols (mnenc_get_symbols) and the streams that
    mnenc_debug_parse finds in the finished file."""
    data = mn.encodeImageFast(img, ctx)
    h, w = img.shape[:2]
    counts, pairs = mn.lastSymbols(ctx, w, h)
    parsed = mn.parseStream(data)
    assert (parsed["lengths"] == counts.reshape(-1)).all()
    return counts, pairs, parsed["streams"]


def scatter_static(counts, pairs, capacity=None, fill=0x5A5A):
    """One call of the new entry point on device copies of counts / pairs; returns (rc, offsets, streams as numpy)."""
    import torch
    torch.cuda.set_device(0)
    L = _lib.load()
    nb = counts.shape[0]
    d_counts = torch.from_numpy(counts.view(np.int16).reshape(-1)).cuda()
    d_pairs = torch.from_numpy(pairs.view(np.int16).reshape(-1)).cuda()
    need = 2 * int(counts.astype(np.int64).sum())
    cap = need if capacity is None else capacity
    d_streams = torch.full((max(cap, 1),), fill, dtype=torch.int16, device="cuda")
    offs = np.full(6 * K + 1, 0xDEAD, np.uint64)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc = L.mnenc_scatter_streams_static(C.c_void_p(d_counts.data_ptr()), C.c_void_p(d_pairs.data_ptr()), K, nb,
                                        C.c_void_p(d_streams.data_ptr()), cap, offs.ctypes.data_as(C.c_void_p), None)
    free_after = torch.cuda.mem_get_info()[0]
    return rc, offs, d_streams.cpu().numpy().view(np.uint16), free_before - free_after


def assert_same_streams(offs, got, want_streams):
    o = offs.astype(np.int64)
    assert len(want_streams) == 6 * K
    for s, want in enumerate(want_streams):
        assert o[s + 1] - o[s] == want.size, f"stream {s}: {o[s + 1] - o[s]} symbols, the encoder's file has {want.size}"
        assert (got[o[s]:o[s + 1]] == want).all(), f"stream {s} differs from the allocating path"


def test_static_scatter_equals_allocating_path_on_consecutive_calls(ctx):
    small = synthetic_rgb(200, 136, 31)    # 25 x 17 blocks, ragged edge: one tile and a bit
    large = synthetic_rgb(640, 512, 32)    # 5120 blocks = 20 tiles: leaves the shared array full of larger offsets
    noise = np.random.default_rng(33).integers(0, 256, (72, 88, 3)).astype(np.uint8)  # deep pursuits: many non-empty streams
    cases = [reference_streams(ctx, im) for im in (small, large, noise)]
    grown = []
    for i in (1, 0, 2, 0, 1):              # large, then small right after it, ... every call sees another call's leftovers
        counts, pairs, want = cases[i]
        rc, offs, got, grew = scatter_static(counts, pairs)
        assert rc == 0, _lib.load().mnenc_last_error()
        assert_same_streams(offs, got, want)
        grown.append(grew)
    # after the first call (module load + address look-up) the entry point must not take GPU memory
    assert all(g <= 0 for g in grown[1:]), grown


def test_static_scatter_empty_and_all_zero(ctx):
    import torch
    L = _lib.load()
    offs = np.full(6 * K + 1, 7, np.uint64)
    assert L.mnenc_scatter_streams_static(None, None, K, 0, None, 0, offs.ctypes.data_as(C.c_void_p), None) == 0
    assert (offs == 0).all()
    black = np.zeros((40, 56, 3), np.uint8)  # every count is 0: all 192 streams are empty, nothing is written
    counts, pairs, want = reference_streams(ctx, black)
    assert counts.max() == 0
    rc, offs, got, _ = scatter_static(counts, pairs, capacity=0)
    assert rc == 0 and (offs == 0).all() and (got == 0x5A5A).all()
    torch.cuda.synchronize()


def test_static_scatter_refuses_what_it_cannot_do(ctx):
    import torch
    L = _lib.load()
    counts, pairs, want = reference_streams(ctx, synthetic_rgb(64, 64, 34))
    offs = np.zeros(6 * K + 1, np.uint64)
    po = offs.ctypes.data_as(C.c_void_p)
    d = torch.zeros(16, dtype=torch.int16, device="cuda")
    p = C.c_void_p(d.data_ptr())
    # beyond the fixed scratch array: refused before anything is launched
    assert L.mnenc_scatter_streams_static(p, p, K, MAX_BLOCKS + 1, p, 16, po, None) == -1
    assert b"1048576" in L.mnenc_last_error()
    assert L.mnenc_scatter_streams_static(p, p, 33, 1, p, 16, po, None) == -1
    # a host pointer must come back as an error, not as an illegal-address fault that poisons the CUDA context
    host = np.zeros(3 * 64, np.uint16)
    assert L.mnenc_scatter_streams_static(host.ctypes.data_as(C.c_void_p), p, K, 64, p, 16, po, None) == -1
    assert b"device memory" in L.mnenc_last_error() or b"pointer" in L.mnenc_last_error()
    # output too small: the sizes are reported, the buffer is left alone
    rc, offs2, got, _ = scatter_static(counts, pairs, capacity=8)
    assert rc == -1 and b"streams_dev holds 8" in L.mnenc_last_error()
    assert int(offs2[-1]) == 2 * int(counts.astype(np.int64).sum()) and (got == 0x5A5A).all()
    # and the GPU is still healthy: the next good call gives the right answer
    rc, offs3, got3, _ = scatter_static(counts, pairs)
    assert rc == 0
    assert_same_streams(offs3, got3, want)
