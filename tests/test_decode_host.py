"""CPU tests of the product's host-side `.mn` reader (imageexperiments_b200/csrc/decode_host.cpp through the host-only
entry points mnenc_stream_info / mnenc_debug_parse): same `lengths` and streams as the reference's readCompressedFast
(golden vectors), as the oracle on larger images, and errors — not crashes — on damaged streams."""
import os
import time

import numpy as np
import pytest

import imageexperiments_b200 as mn

FIXTURE = "/root/reference/Data/r0c1de5e1t_3_5.mn"


@pytest.fixture(scope="module")
def gd():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_decode.npz"))


def same_parse(p, lengths, streams):
    assert (p["lengths"] == lengths).all()
    assert len(p["streams"]) == len(streams)
    for i, (a, b) in enumerate(zip(p["streams"], streams)):
        assert a.shape == b.shape and (a == b).all(), i


def test_parse_golden(gd):
    for name in gd["names"]:
        data = gd[f"mn_{name}"].tobytes()
        p = mn.parseStream(data)
        sizes = gd[f"sizes_{name}"]
        offs = np.concatenate([[0], np.cumsum(sizes)])
        same_parse(p, gd[f"len_{name}"], [gd[f"sym_{name}"][offs[i]:offs[i + 1]] for i in range(len(sizes))])
        h, w = gd[f"rgb_{name}"].shape[:2]
        assert (p["width"], p["height"], p["blockSize"]) == (w, h, 8) and p["K"] == len(sizes) // 6


def test_stream_info_is_the_header(gd, golden):
    info = mn.streamInfo(gd["mn_coarse"].tobytes())
    assert (info["Quant"] == mn.createQuantizationTables(32, 8, 1.0)).all()
    hdr = golden["fixture_header"].tobytes()  # the reference's own fixture: 4928 x 3264, K 32, tables of quality 3.5
    info = mn.streamInfo(hdr + b"\0" * 64)
    assert (info["width"], info["height"], info["K"], info["blockSize"]) == (4928, 3264, 32, 8)
    assert (info["Quant"] == mn.createQuantizationTables(32, 8, 3.5)).all()


def test_parse_vs_oracle_larger_images(oraclelib, oracle_ctx_factory):
    from imageexperiments_b200.synth import synthetic_rgb
    rng = np.random.default_rng(5)
    cases = [(synthetic_rgb(512, 384, 41), 3.5), (synthetic_rgb(333, 257, 42), 8.0), (synthetic_rgb(640, 120, 43), 1.0),
             (np.full((1600, 1608, 3), 200, np.uint8), 3.5),                       # runs over the 32 769 chunk limit
             (rng.integers(0, 256, (96, 160, 3)).astype(np.uint8), 8.0)]         # wide alphabets, long codes
    for img, q in cases:
        oc = oracle_ctx_factory(mn.createQuantizationTables(32, 8, q))
        data = oc.encode(img)
        oc.close()
        want = oraclelib.parse(data)
        same_parse(mn.parseStream(data), want["lengths"], want["streams"])


def test_damaged_streams_give_errors(gd):
    data = gd["mn_synth"].tobytes()
    bad = [b"", data[:10], b"\0\0\0\0" + data[4:], data[: len(data) // 2], data[:300],
           data[:12] + bytes([33]) + data[13:],      # K = 33
           data[:13] + bytes([4]) + data[14:],       # block size 4
           data[:4] + b"\x7f\xff\xff\xff" + data[8:]]  # absurd width
    for b in bad:
        with pytest.raises(mn.MnencError) as e:
            mn.parseStream(b)
        assert e.value.code == -5, len(b)
    rng = np.random.default_rng(9)
    arr = np.frombuffer(data, np.uint8)
    for _ in range(300):  # random damage after the header: either rejected or parsed, never a crash
        a = arr.copy()
        pos = rng.integers(206, a.size, 4)
        a[pos] ^= rng.integers(1, 256, 4).astype(np.uint8)
        try:
            mn.parseStream(a.tobytes())
        except mn.MnencError as e:
            assert e.code == -5


@pytest.mark.skipif(not os.path.exists(FIXTURE), reason="reference tree absent (GPU box)")
def test_parse_reference_fixture(oraclelib):
    with open(FIXTURE, "rb") as f:
        data = f.read()
    t = time.perf_counter()
    p = mn.parseStream(data)
    ms = (time.perf_counter() - t) * 1e3
    want = oraclelib.parse(data)
    same_parse(p, want["lengths"], want["streams"])
    same_parse(mn.parseStream(data, threads=8), want["lengths"], want["streams"])  # parallel reader
    print(f"fixture: {len(data)} B, {sum(len(s) for s in p['streams'])} symbols parsed (twice) in {ms:.1f} ms")


def crafted(oraclelib, K, w, h, counts, pairs):
    """counts [nb,3], pairs [nb,3,K,2] (raw symbols, any u16 values) -> stream written by the oracle's writeCompressed."""
    return oraclelib.write_file(K, 8, w, h, np.arange(1, 3 * K + 1, dtype=np.float32), counts, pairs)


def expected_streams(K, counts, pairs):
    """Stream (layer, depth, part) = entries of the blocks with count > depth, block order (CompressedImage.cpp:612-629)."""
    out = []
    for ch in range(3):
        for d in range(K):
            sel = counts[:, ch] > d
            for part in range(2):
                out.append(pairs[sel, ch, d, part])
    return out


@pytest.mark.parametrize("kind", ["uniform16", "geometric", "constant", "sparse", "bimodal", "ramp"])
@pytest.mark.parametrize("K", [1, 8, 32])
def test_parse_crafted_symbol_distributions(oraclelib, kind, K):
    """Symbol statistics the matching pursuit never produces, to reach every branch of the reader: 16-bit alphabets with
    codes far beyond the 12-bit table, raw vs Elias-Fano symbol lists, Golomb-coded sequences of every parameter, runs across
    the 32 769 chunk limit, single-symbol and empty streams."""
    rng = np.random.default_rng(sum(map(ord, kind)) * 100 + K)
    w, h = 8 * 70, 8 * 64
    nb = 70 * 64
    if kind == "constant":
        w, h, nb = 8 * 300, 8 * 120, 36000  # one run of 36 000 > 32 769
    if kind == "sparse":
        counts = (rng.random((nb, 3)) < 0.002).astype(np.uint16) * K  # nearly empty streams
    elif kind == "constant":
        counts = np.full((nb, 3), K, np.uint16)
    else:
        counts = rng.integers(0, K + 1, (nb, 3)).astype(np.uint16)
    shape = (nb, 3, K, 2)
    if kind == "uniform16":
        pairs = rng.integers(0, 65536, shape)
    elif kind == "geometric":
        pairs = np.minimum(rng.geometric(0.02, shape) - 1, 65535)
    elif kind == "constant":
        pairs = np.full(shape, 4321)
        pairs[:, 1] = 0          # leading zeros: not a repeat of the initial prevVal (Huffman.cpp:253-255)
        pairs[::7000, 2, 0, 0] = 9  # a few breaks
    elif kind == "sparse":
        pairs = rng.integers(0, 3, shape)
    elif kind == "bimodal":
        pairs = np.where(rng.random(shape) < 0.97, rng.integers(0, 4, shape), rng.integers(0, 65536, shape))  # short + very long codes
    else:
        pairs = (np.arange(np.prod(shape)).reshape(shape) // 3) % 50000
    pairs = pairs.astype(np.uint16)
    pairs[:, :, 0, 1] %= 16384  # DC chain: zig-zag of a difference must fit 16 bits to survive (CompressedImage.cpp:487-505 truncates)
    pairs[np.broadcast_to(np.arange(K)[None, None, :, None] >= counts[:, :, None, None], shape)] = 0  # entries beyond count are not coded
    data = crafted(oraclelib, K, w, h, counts, pairs)
    p = mn.parseStream(data)
    assert p["K"] == K and (p["width"], p["height"]) == (w, h)
    same_parse(p, counts.reshape(-1), expected_streams(K, counts, pairs))
    op = oraclelib.parse(data)  # and the oracle's own reader agrees
    same_parse(p, op["lengths"], op["streams"])


@pytest.mark.parametrize("kind", ["uniform8", "uniform16", "bimodal", "skewed", "constant"])
def test_parallel_reader_equals_serial(oraclelib, kind):
    """The speculative multi-threaded Huffman reader (decode_host.cpp, MNENC_DECODE_THREADS) must return exactly what the serial
    reader returns: payloads that fall into step quickly (skewed), late (bimodal: long codes) or never (fixed-length-like
    codes of a uniform alphabet: every chunk is walked serially), and sequences too short or Golomb coded (left serial)."""
    rng = np.random.default_rng(sum(map(ord, kind)))
    K, bw, bh = 4, 300, 220
    nb = bw * bh
    counts = rng.integers(0, K + 1, (nb, 3)).astype(np.uint16) if kind != "constant" else np.full((nb, 3), K, np.uint16)
    shape = (nb, 3, K, 2)
    if kind == "uniform8":
        pairs = rng.integers(0, 256, shape)
    elif kind == "uniform16":
        pairs = rng.integers(0, 65536, shape)
    elif kind == "bimodal":
        pairs = np.where(rng.random(shape) < 0.9, rng.integers(0, 3, shape), rng.integers(0, 40000, shape))
    elif kind == "skewed":
        pairs = np.minimum((rng.pareto(1.2, shape) * 3).astype(np.int64), 65535)
    else:
        pairs = np.full(shape, 77)
    pairs = pairs.astype(np.uint16)
    pairs[:, :, 0, 1] %= 16384
    pairs[np.broadcast_to(np.arange(K)[None, None, :, None] >= counts[:, :, None, None], shape)] = 0
    data = crafted(oraclelib, K, 8 * bw, 8 * bh, counts, pairs)
    serial = mn.parseStream(data)
    same_parse(serial, counts.reshape(-1), expected_streams(K, counts, pairs))
    for threads in (2, 5):
        par = mn.parseStream(data, threads=threads)
        same_parse(par, serial["lengths"], serial["streams"])


def test_parallel_reader_on_damaged_streams(gd, oracle_ctx_factory):
    """Damage inside large payloads: the parallel reader reports an error or the serial reader's symbols, never anything else."""
    from imageexperiments_b200.synth import synthetic_rgb
    oc = oracle_ctx_factory(mn.createQuantizationTables(32, 8, 3.5))
    data = np.frombuffer(oc.encode(synthetic_rgb(1400, 1000, 3)), np.uint8)
    oc.close()
    rng = np.random.default_rng(4)
    same_parse(mn.parseStream(data.tobytes(), threads=4), *(lambda p: (p["lengths"], p["streams"]))(mn.parseStream(data.tobytes())))
    for _ in range(40):
        a = data.copy()
        pos = rng.integers(206, a.size, 2)
        a[pos] ^= rng.integers(1, 256, 2).astype(np.uint8)
        try:
            want = mn.parseStream(a.tobytes())
        except mn.MnencError:
            want = None
        try:
            got = mn.parseStream(a.tobytes(), threads=4)
        except mn.MnencError as e:
            assert e.code == -5
            got = None
        assert (want is None) == (got is None)
        if want is not None:
            same_parse(got, want["lengths"], want["streams"])
