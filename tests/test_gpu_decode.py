"""GPU parity tests of the decoder: mnenc_decode_rgb8* must give the pixels of the reference's
decodeImageFast, bit for bit — checked against golden vectors made by the reference's decoder, against the CPU oracle
(itself pinned on the reference) for streams from the GPU encoder, and on BASELINE.json's full-size geometry."""
import os

import numpy as np
import pytest

import imageexperiments_b200 as mn
from imageexperiments_b200.synth import synthetic_rgb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gd():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_decode.npz"))


@pytest.fixture(scope="module")
def ctx(dictionary):
    c = mn.createCompressionContextFast(32, 8, 3.5, dictionary=dictionary, device=0)
    yield c
    c.close()


def test_decode_golden(ctx, gd):
    """Streams written by the reference encoder (K 32 and 8, all qualities, ragged / empty / run-length coded streams)."""
    for name in gd["names"]:
        got = mn.decodeImageFast(gd[f"mn_{name}"].tobytes(), ctx)
        want = gd[f"rgb_{name}"]
        assert got.shape == want.shape and (got == want).all(), name


@pytest.mark.parametrize("w,h,q,seed", [(512, 512, 3.5, 5), (200, 136, 3.5, 4), (1023, 77, 6.0, 11), (8, 8, 3.5, 1), (24, 40, 1.0, 2),
                                        (333, 257, 8.0, 7), (2048, 16, 3.5, 8), (16, 2048, 3.5, 9)])
def test_round_trip_equals_oracle_decoder(ctx, oracle_ctx_factory, w, h, q, seed):
    img = synthetic_rgb(w, h, seed)
    quant = mn.createQuantizationTables(32, 8, q)
    ctx.setQuant(quant)
    data = mn.encodeImageFast(img, ctx)
    ctx.setQuant(mn.createQuantizationTables(32, 8, 3.5))  # the decoder must take the tables from the stream
    got = mn.decodeImageFast(data, ctx)
    oc = oracle_ctx_factory(quant)
    want = oc.decode(data)
    oc.close()
    assert got.shape == img.shape and (got == want).all()
    assert np.abs(got.astype(int) - img).mean() < 20  # and it is the picture


def test_decode_edge_images(ctx, oracle_ctx_factory):
    rng = np.random.default_rng(3)
    quant = mn.createQuantizationTables(32, 8, 3.5)
    ctx.setQuant(quant)
    oc = oracle_ctx_factory(quant)
    for img in (np.zeros((40, 56, 3), np.uint8), np.full((1600, 1608, 3), 77, np.uint8),
                rng.integers(0, 256, (72, 88, 3)).astype(np.uint8), np.full((9, 9, 3), 255, np.uint8)):
        data = mn.encodeImageFast(img, ctx)
        assert (mn.decodeImageFast(data, ctx) == oc.decode(data)).all(), img.shape
    oc.close()


def test_decode_max_quality(ctx, oracle_ctx_factory):
    """All-ones tables (Compression.cpp:104-114): 32 picks in most blocks, deep detail-group walks."""
    img = synthetic_rgb(96, 64, 31)
    ctx.setQuant(np.ones(96, np.float32))
    data = mn.encodeImageFast(img, ctx)
    oc = oracle_ctx_factory(np.ones(96, np.float32))
    want = oc.decode(data)
    oc.close()
    ctx.setQuant(mn.createQuantizationTables(32, 8, 3.5))
    got = mn.decodeImageFast(data, ctx)
    assert (got == want).all()
    assert np.abs(got.astype(int) - img).max() <= 8


def test_decode_into_device_tensor_and_zero_copy(ctx):
    import torch
    img = synthetic_rgb(640, 360, 17)
    data = mn.encodeImageFast(img, ctx, zero_copy=True)   # MnBuffer straight back into the decoder
    want = mn.decodeImageFast(data, ctx)
    out = torch.zeros((360, 640, 3), dtype=torch.uint8, device="cuda:0")
    assert mn.decodeImageFast(data, ctx, out=out) is out
    assert (out.cpu().numpy() == want).all()
    view = mn.decodeImageFast(bytes(data), ctx, zero_copy=True)
    assert (view == want).all()
    with pytest.raises(ValueError):
        mn.decodeImageFast(data, ctx, out=torch.zeros((360, 648, 3), dtype=torch.uint8, device="cuda:0"))
    with pytest.raises(mn.MnencError) as e:  # too small for the image in the stream
        mn.decodeImageFast(data, ctx, out=torch.zeros((180, 640, 3), dtype=torch.uint8, device="cuda:0"))
    assert e.value.code == -1


def test_batch_decode_equals_one_at_a_time(ctx, gd):
    """mnenc_decode_batch_rgb8: mixed sizes, K and qualities over the lanes; each image as from decodeImageFast."""
    names = list(gd["names"]) * 3
    streams = [gd[f"mn_{n}"].tobytes() for n in names]
    imgs = [synthetic_rgb(640, 360, 100 + i) for i in range(5)]
    streams += mn.encodeImageBatch(imgs, ctx, zero_copy=True)
    got = mn.decodeImageBatch(streams, ctx)
    for n, g in zip(names, got):
        assert (g == gd[f"rgb_{n}"]).all(), n
    for s, g in zip(streams[len(names):], got[len(names):]):
        assert (g == mn.decodeImageFast(s, ctx)).all()
    assert mn.decodeImageBatch([], ctx) == []
    with pytest.raises(mn.MnencError) as e:  # one damaged stream fails the batch
        mn.decodeImageBatch(streams[:4] + [streams[0][:50]], ctx)
    assert e.value.code == -5
    views = mn.decodeImageBatch(streams[:3], ctx, zero_copy=True)
    assert all((v == gd[f"rgb_{n}"]).all() for v, n in zip(views, names))


def test_damaged_streams_are_rejected(ctx, gd):
    data = gd["mn_synth"].tobytes()
    for bad in (b"\0\0\0\0" + data[4:], data[: len(data) // 2], data[:100]):
        with pytest.raises(mn.MnencError) as e:
            mn.decodeImageFast(bad, ctx)
        assert e.value.code == -5
    assert (mn.decodeImageFast(data, ctx) == gd["rgb_synth"]).all()  # the context is still usable


def test_encode_after_decode_is_unchanged(ctx, gd):
    """The decoder shares device buffers with the encoder; interleaving them must not change either result."""
    img = synthetic_rgb(264, 200, 23)
    a = mn.encodeImageFast(img, ctx)
    mn.decodeImageFast(gd["mn_gray"].tobytes(), ctx)
    b = mn.encodeImageFast(img, ctx)
    assert a == b


def test_full_size_round_trip(ctx, oracle_ctx_factory, reflib):
    """BASELINE.json configs[1] geometry (4928 x 3264): pixels == oracle decoder; the reference's own PSNR of the result."""
    img = synthetic_rgb(4928, 3264, 12345)
    quant = mn.createQuantizationTables(32, 8, 3.5)
    ctx.setQuant(quant)
    data = mn.encodeImageFast(img, ctx)
    got = mn.decodeImageFast(data, ctx)
    oc = oracle_ctx_factory(quant)
    want = oc.decode(data)
    oc.close()
    assert (got == want).all()
    assert reflib.psnr(img, got) > 40.0
