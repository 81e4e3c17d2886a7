"""Pins the decoder half of the CPU oracle (oracle/mn_oracle.cpp: mno_parse, mno_decode) against golden
vectors made by the reference's own readCompressedFast / decodeImageFast (tests/golden/make_golden_decode.py), against the
reference build itself where oracle/_ref is present, and against the reference's fixture Data/r0c1de5e1t_3_5.mn where the
reference tree is present (this container only)."""
import hashlib
import os

import numpy as np
import pytest

FIXTURE = "/root/reference/Data/r0c1de5e1t_3_5.mn"


@pytest.fixture(scope="module")
def gd():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_decode.npz"))


def test_parse_golden(oraclelib, gd):
    for name in gd["names"]:
        p = oraclelib.parse(gd[f"mn_{name}"].tobytes())
        assert (p["lengths"] == gd[f"len_{name}"]).all(), name
        assert [len(s) for s in p["streams"]] == list(gd[f"sizes_{name}"]), name
        flat = np.concatenate(p["streams"]) if p["streams"] else np.zeros(0, np.uint16)
        assert (flat == gd[f"sym_{name}"]).all(), name


def test_decode_golden(oracle_ctx_factory, gd):
    oc = oracle_ctx_factory(np.ones(96, np.float32))  # the tables come from the file, not from the context
    for name in gd["names"]:
        got = oc.decode(gd[f"mn_{name}"].tobytes())
        want = gd[f"rgb_{name}"]
        assert got.shape == want.shape and (got == want).all(), name
    oc.close()


def test_decode_vs_reference(oracle_ctx_factory, reflib):
    from imageexperiments_b200.synth import synthetic_rgb
    oc = oracle_ctx_factory(np.ones(96, np.float32))
    for (w, h, q, seed) in [(96, 64, 3.5, 31), (129, 71, 6.0, 32), (40, 200, 2.0, 33)]:
        ctx = reflib.create(32, 8, q)
        data = reflib.encode(ctx, synthetic_rgb(w, h, seed))
        reflib.destroy(ctx)
        assert (oc.decode(data) == reflib.decode(data)).all(), (w, h, q)
        rp, op = reflib.parse(data), oc.lib.parse(data)
        assert (rp["lengths"] == op["lengths"]).all()
        assert all((a == b).all() for a, b in zip(rp["streams"], op["streams"]))
    oc.close()


def test_invalid_streams_are_rejected(oraclelib, gd):
    data = gd["mn_synth"].tobytes()
    with pytest.raises(ValueError):
        oraclelib.parse(b"\0\0\0\0" + data[4:])      # magic, CompressedImage.cpp:721-723
    with pytest.raises(ValueError):
        oraclelib.parse(data[: len(data) // 2])       # truncated: Huffman payload without its pseudo-EOF, Huffman.cpp:243


@pytest.mark.skipif(not os.path.exists(FIXTURE), reason="reference tree absent (GPU box)")
def test_decode_reference_fixture(oracle_ctx_factory, gd):
    """The reference's own 16 MP stream (SURVEY.md §8c pin 2): same pixels as the reference's decoder."""
    oc = oracle_ctx_factory(np.ones(96, np.float32))
    with open(FIXTURE, "rb") as f:
        img = oc.decode(f.read())
    oc.close()
    assert tuple(img.shape[:2]) == tuple(gd["fixture_dims"])
    assert hashlib.sha256(img.tobytes()).hexdigest() == gd["fixture_sha256"].tobytes().decode()
