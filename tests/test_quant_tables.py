"""createQuantizationTables (float overload, CompressedImage.cpp:168-210): golden vectors from the reference build
and the header of the reference's own fixture Data/r0c1de5e1t_3_5.mn (known answer at quality 3.5)."""
import numpy as np

import imageexperiments_b200 as mn


def test_against_golden(golden):
    for q, want in zip(golden["quant_q"], golden["quant_tables"]):
        got = mn.createQuantizationTables(32, 8, float(q))
        assert (got == want).all(), q


def test_fixture_header_known_answer(golden):
    hdr = golden["fixture_header"].tobytes()
    assert hdr[:4] == b"MN24"
    assert int.from_bytes(hdr[4:8], "big") == 4928 and int.from_bytes(hdr[8:12], "big") == 3264
    assert hdr[12] == 32 and hdr[13] == 8
    quant = np.frombuffer(hdr[14:206], ">u2").astype(np.float32)
    assert (mn.createQuantizationTables(32, 8, 3.5) == quant).all()
    assert quant[:8].tolist() == [8, 1, 1, 3, 3, 5, 9, 8]


def test_against_reference_sweep(reflib):
    for q in np.arange(0.25, 9.01, 0.25):
        assert (mn.createQuantizationTables(32, 8, float(q)) == reflib.quant_tables(32, 8, float(q))).all(), q
    for K in (1, 8, 16):
        assert (mn.createQuantizationTables(K, 8, 3.5) == reflib.quant_tables(K, 8, 3.5)).all(), K


def test_calculate_psnr_equals_reference(reflib):
    """compressed::calculatePSNR (CompressedImage.cpp:343-357), incl. its 3 * 255 peak."""
    import imageexperiments_b200 as mn
    from imageexperiments_b200.synth import synthetic_rgb
    a = synthetic_rgb(96, 64, 1)
    rng = np.random.default_rng(0)
    b = np.clip(a.astype(int) + rng.integers(-6, 7, a.shape), 0, 255).astype(np.uint8)
    assert abs(mn.calculatePSNR(a, b) - reflib.psnr(a, b)) < 1e-9
    assert mn.calculatePSNR(a, a) == float("inf")
