"""Pins the CPU oracle (oracle/mn_oracle.cpp): against golden vectors generated from the reference build, and — when
oracle/_ref is present — against the unmodified reference itself. The entropy cases restate the reference's own
GoogleTest cases (Testing/BitBufferTests.cpp, Testing/HuffmanTest.cpp): empty input, all integer-partition frequency
profiles, all-ones symbol vs pseudo-EOF, wide alphabets, long runs across the 0x8000 RLE chunk limit."""
import numpy as np
import pytest

from conftest import bits_of


def test_entropy_golden(oraclelib, golden):
    for i in range(int(golden["ent_count"])):
        data = golden[f"ent_{i}_in"]
        blob, nbits = oraclelib.huffman_or_golomb(data)
        assert nbits == int(golden[f"ent_{i}_nbits"]), i
        assert (bits_of(blob, nbits) == bits_of(golden[f"ent_{i}_bits"].tobytes(), nbits)).all(), i
        assert (oraclelib.rle(data) == golden[f"ent_{i}_rle"]).all(), i


def test_image_golden(oracle_ctx_factory, golden):
    import imageexperiments_b200 as mn
    for name in golden["img_names"]:
        img = golden[f"img_{name}"]
        oc = oracle_ctx_factory(mn.createQuantizationTables(32, 8, float(golden[f"q_{name}"])))
        counts, pairs = oc.symbols(img)
        assert (counts == golden[f"sym_{name}_counts"]).all(), name
        assert (pairs == golden[f"sym_{name}_pairs"]).all(), name
        assert oc.encode(img) == golden[f"file_{name}"].tobytes(), name
        oc.close()


def test_zigzag_golomb_eliasfano_vs_reference(oraclelib, reflib):
    rng = np.random.default_rng(5)
    for x in [0, 1, -1, 2, -2, 32767, -32768, 100000, -100000, 2**31 - 1, -2**31]:
        assert oraclelib.zigzag(x) == reflib.zigzag(x)
    for m in list(range(1, 256)) + [511, 512, 1023, 1024, 2047]:  # BitBufferTests.cpp:169-182
        vals = np.concatenate([np.arange(0, 40), rng.integers(0, 5000, 20)]).astype(np.uint32)
        for v in vals[:12]:
            assert oraclelib.golomb_length(int(v), m) == reflib.golomb_length(int(v), m)
        a, na = oraclelib.golomb(vals, m)
        b, nb = reflib.golomb(vals, m)
        assert na == nb and a == b, m
        assert na == sum(reflib.golomb_length(int(v), m) for v in vals)
    for n, mx in [(1, 1), (5, 255), (100, 4095), (1000, 65535), (17, 1023)]:  # BitBufferTests.cpp:197-247
        seq = np.sort(rng.integers(0, mx + 1, n)).astype(np.uint16)
        assert oraclelib.elias_fano_bound(n, mx) == reflib.elias_fano_bound(n, mx)
        a, na = oraclelib.elias_fano(seq, mx)
        b, nb = reflib.elias_fano(seq, mx)
        assert na == nb and a == b
        assert na <= reflib.elias_fano_bound(n, mx)


def _partitions(n, maxpart=None):
    maxpart = maxpart or n
    if n == 0:
        yield []
        return
    for k in range(min(n, maxpart), 0, -1):
        for rest in _partitions(n - k, k):
            yield [k] + rest


def test_huffman_vs_reference(oraclelib, reflib):
    rng = np.random.default_rng(6)
    cases = [np.zeros(0, np.uint16)]
    for part in _partitions(10):  # HuffmanTest.cpp:16-49
        cases.append(rng.permutation(np.concatenate([np.full(c, 11 * i, np.uint16) for i, c in enumerate(part)])))
    cases.append(np.full(7, 1, np.uint16))
    cases.append(np.full(5, 0xFFFF, np.uint16))
    cases.append(rng.integers(0, 65536, 100000).astype(np.uint16))  # HuffmanTest.cpp:93-111
    for _ in range(60):
        n = int(rng.integers(1, 4000))
        kind = int(rng.integers(0, 4))
        d = [rng.integers(0, 3, n), rng.geometric(0.02, n) - 1, rng.integers(0, 3000, n), np.minimum(rng.pareto(0.8, n) * 30, 65535)][kind]
        cases.append(np.asarray(d).astype(np.uint16))
    for d in cases:
        a, na = oraclelib.huffman(d)
        b, nb = reflib.huffman(d)
        assert na == nb and a == b, len(d)
        assert (reflib.huffman_decode(a) == d).all()
        a, na = oraclelib.huffman_or_golomb(d)
        b, nb = reflib.huffman_or_golomb(d)
        assert na == nb and a == b, len(d)


@pytest.mark.parametrize("runs", [[(5, 4), (7, 1)], [(9, 32770)], [(0, 3), (1, 1), (1, 1)], [(3, 50000)], [(2, 32769), (2, 1)],
                                  [(0, 1)], [(4, 32768)], [(4, 65538), (5, 2)], [(8, 98307)]])
def test_rle_vs_reference(oraclelib, reflib, runs):  # HuffmanTest.cpp:126-178
    d = np.concatenate([np.full(n, v, np.uint16) for v, n in runs])
    got = oraclelib.rle(d)
    assert (got == reflib.rle(d)).all()
    assert (reflib.rle_decode(got) == d).all()


def test_rle_random_vs_reference(oraclelib, reflib):
    rng = np.random.default_rng(7)
    for _ in range(50):
        n = int(rng.integers(0, 3000))
        d = np.repeat(rng.integers(0, 4, n), rng.integers(1, 6, n)).astype(np.uint16)
        assert (oraclelib.rle(d) == reflib.rle(d)).all()


def test_matching_pursuit_vs_reference(oracle_ctx_factory, reflib):
    from imageexperiments_b200.synth import synthetic_rgb
    for (w, h, q, seed) in [(96, 64, 3.5, 11), (61, 45, 3.5, 12), (48, 40, 8.0, 13), (64, 64, 1.0, 14)]:
        ctx = reflib.create(32, 8, q)
        quant = reflib.get_quant(ctx)
        oc = oracle_ctx_factory(quant)
        img = synthetic_rgb(w, h, seed)
        rc, rp = reflib.symbols(ctx, img)
        oc_c, oc_p = oc.symbols(img)
        assert (rc == oc_c).all() and (rp == oc_p).all(), (w, h, q)
        ref_bytes = reflib.encode(ctx, img)
        assert oc.encode(img) == ref_bytes
        parsed = reflib.parse(ref_bytes)  # the reference's own reader recovers the same symbols
        assert (parsed["lengths"].reshape(-1, 3) == rc).all()
        oc.close()
        reflib.destroy(ctx)


def test_max_quality_vs_reference(oracle_ctx_factory, reflib):
    from imageexperiments_b200.synth import synthetic_rgb
    ctx = reflib.create(32, 8, 3.5)
    reflib.set_quant(ctx, np.ones(96, np.float32))  # Compression.cpp:104-114 "max"
    oc = oracle_ctx_factory(np.ones(96, np.float32))
    img = synthetic_rgb(24, 16, 3)
    assert oc.encode(img) == reflib.encode(ctx, img)
    oc.close()
    reflib.destroy(ctx)


def test_dictionary_via_callbacks_equals_direct_export(reflib, dictionary):
    """The drop-in shim recovers the dictionary through the opaque Dynamic callbacks; both routes
    must give the same floats (CompressedImage.cpp:252-289,331-339)."""
    ctx = reflib.create(32, 8, 3.5)
    assert (reflib.dynamic_probe(ctx, 0, -1) == dictionary.base).all()
    offs = np.concatenate([[0], np.cumsum(dictionary.detail_rows)])
    for ch in range(3):
        for atom in (0, 1, 77, 508, 509):
            full = reflib.dynamic_probe(ctx, ch, atom)
            assert full.shape[0] == 510 + dictionary.detail_rows[atom]
            assert (full[510:] == dictionary.detail[ch][offs[atom]:offs[atom + 1]]).all()
    reflib.destroy(ctx)
