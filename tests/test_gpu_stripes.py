"""GPU tests of the stripe protocol (BASELINE.json configs[3]): ONE image sharded by block columns over
`world` encoder contexts must give the very bytes of the single-GPU encode == the oracle's == the reference's.

On a one-GPU box the ranks are threads of this process (ThreadComm: every rank has its own context and buffers on
cuda:0, the exchanges are the same calls); with >= 2 GPUs the same protocol also runs as one process per GPU over NCCL."""
import os
import socket
import subprocess
import sys
import threading

import numpy as np
import pytest

import imageexperiments_b200 as mn
from imageexperiments_b200.synth import synthetic_rgb

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MAX_WORLD = 5


@pytest.fixture(scope="module")
def ctxs(dictionary):
    cs = [mn.createCompressionContextFast(32, 8, 3.5, dictionary=dictionary, device=0) for _ in range(MAX_WORLD)]
    yield cs
    for c in cs:
        c.close()


def run_stripes(img, world, ctxs, quant=None, device_stripes=False):
    """Encode `img` over `world` thread-ranks; returns (bytes from rank 0, per-rank (counts, pairs))."""
    import torch
    comms = mn.ThreadComm.make(world)
    results, errors, symbols = [None] * world, [], [None] * world
    h, w = img.shape[:2]
    cols = mn.stripeColumns((w + 7) // 8, world)

    def work(r):
        try:
            if quant is not None:
                ctxs[r].setQuant(quant)
            if device_stripes:
                x0, x1 = 8 * cols[r][0], min(8 * cols[r][1], w)
                t = torch.from_numpy(np.ascontiguousarray(img[:, x0:x1])).cuda()
                results[r] = mn.encodeImageStripes(t, ctxs[r], comms[r], width=w)
            else:
                results[r] = mn.encodeImageStripes(img, ctxs[r], comms[r])
            ws = min(8 * cols[r][1], w) - 8 * cols[r][0]
            if ws > 0:
                symbols[r] = mn.lastSymbols(ctxs[r], ws, h)
        except BaseException as e:  # noqa: BLE001 - make the other ranks fail instead of waiting forever
            errors.append((r, e))
            comms[r].s.barrier.abort()

    ts = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(300)
    real = [e for e in errors if not isinstance(e[1], threading.BrokenBarrierError)]
    assert not errors, real or errors
    assert all(results[r] is None for r in range(1, world))
    return results[0], symbols


@pytest.mark.parametrize("w,h,seed", [(512, 512, 5), (200, 136, 4), (1023, 77, 11), (8, 8, 1), (24, 40, 2)])
@pytest.mark.parametrize("world", [1, 2, 3, 5])
def test_stripes_equal_single_gpu_and_oracle(ctxs, oracle_ctx_factory, w, h, seed, world):
    img = synthetic_rgb(w, h, seed)
    q = mn.createQuantizationTables(32, 8, 3.5)
    got, symbols = run_stripes(img, world, ctxs, q)
    oc = oracle_ctx_factory(q)
    assert got == oc.encode(img)
    # the ranks' symbols are the oracle's symbols of their block columns
    cols = mn.stripeColumns((w + 7) // 8, world)
    for r, (bx0, bx1) in enumerate(cols):
        if bx1 > bx0:
            oc_c, oc_p = oc.symbols(img, bx0, bx1)
            assert (symbols[r][0] == oc_c).all() and (symbols[r][1] == oc_p).all(), r
    oc.close()


@pytest.mark.parametrize("world", [2, 4])
def test_stripes_edge_images(ctxs, oracle_ctx_factory, world):
    q = mn.createQuantizationTables(32, 8, 3.5)
    oc = oracle_ctx_factory(q)
    black = np.zeros((40, 56, 3), np.uint8)          # every stream empty on every rank
    gray = np.full((1600, 1608, 3), 77, np.uint8)    # 40 200 identical blocks: one run per stream over all ranks, 32 769 chunk rule
    half = np.zeros((64, 256, 3), np.uint8)          # left ranks: empty streams; DC chain starts on a later rank
    half[:, 160:] = synthetic_rgb(96, 64, 3)
    rng = np.random.default_rng(3)
    noise = rng.integers(0, 256, (72, 88, 3)).astype(np.uint8)
    for img in (black, gray, half, noise):
        got, _ = run_stripes(img, world, ctxs, q)
        assert got == oc.encode(img), img.shape
    oc.close()


def test_stripes_quality_sweep_and_device_input(ctxs, oracle_ctx_factory):
    img = synthetic_rgb(256, 256, 6)
    for qual in (1.0, 8.0, "max"):
        q = np.ones(96, np.float32) if qual == "max" else mn.createQuantizationTables(32, 8, qual)
        oc = oracle_ctx_factory(q)
        got, _ = run_stripes(img, 3, ctxs, q, device_stripes=(qual == 8.0))
        assert got == oc.encode(img), qual
        oc.close()


def test_stripes_full_size_equal_single_gpu(ctxs, reflib):
    """BASELINE.json configs[1] geometry over 4 ranks: same bytes as the single-GPU encode, and the unmodified
    reference decoder round-trips them."""
    w, h = 4928, 3264
    img = synthetic_rgb(w, h, 12345)
    q = mn.createQuantizationTables(32, 8, 3.5)
    ctxs[0].setQuant(q)
    want = mn.encodeImageFast(img, ctxs[0])
    got, _ = run_stripes(img, 4, ctxs, q)
    assert got == want
    dec = reflib.decode(got)
    assert dec.shape == img.shape and reflib.psnr(img, dec) > 40.0


def test_stripe_protocol_order_is_checked(ctxs):
    import ctypes as C
    from imageexperiments_b200 import _lib
    out, n = C.c_void_p(), C.c_size_t()
    with pytest.raises(mn.MnencError) as e:
        _lib.check(_lib.load().mnenc_stripe_finish(ctxs[0]._h, C.byref(out), C.byref(n)))
    assert e.value.code == -3


def test_stripes_nccl_two_processes(dictionary):
    """One process per GPU over NCCL (needs >= 2 GPUs; the thread-rank tests above cover the protocol on one GPU)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "tools", "stripe_check.py")], capture_output=True, text=True,
                       timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "stripe_check OK" in r.stdout


def test_stripes_other_iteration_budget(dictionary, oracle_ctx_factory):
    """K = 8: 48 symbol streams instead of 192; the records, tables and exchanges of the protocol all scale with K."""
    K = 8
    cs = [mn.createCompressionContextFast(K, 8, 3.5, dictionary=dictionary, device=0) for _ in range(3)]
    try:
        img = synthetic_rgb(264, 96, 77)
        got, _ = run_stripes(img, 3, cs)
        oc = oracle_ctx_factory(cs[0].Quant, K)
        assert got == oc.encode(img)
        oc.close()
    finally:
        for c in cs:
            c.close()


def test_in_library_exchange_single_rank(dictionary, oracle_ctx_factory):
    """mnenc_encode_striped_rgb8 over a one-rank NCCL communicator: the in-library exchanges (sparse histogram gather + merge,
    dealt code tables, segment reduce) must reproduce the plain encode byte for byte - host, device and stripe-only inputs."""
    import torch
    ctx = mn.createCompressionContextFast(32, 8, 3.5, dictionary=dictionary, device=0)
    try:
        assert mn.initStripeComm(ctx) == (0, 1)
        oc = oracle_ctx_factory(ctx.Quant)
        for (w, h, seed) in ((512, 512, 1), (1023, 77, 2), (8, 8, 3), (5, 3, 4)):
            img = synthetic_rgb(w, h, seed)
            want = oc.encode(img)
            tm = {}
            assert mn.encodeImageStriped(img, ctx, timings=tm) == want, (w, h)
            assert set(tm) == set(mn.STRIPE_PHASES)
            assert mn.encodeImageStriped(torch.from_numpy(img).cuda(), ctx, width=w) == want, (w, h)
            assert mn.encodeImageFast(img, ctx) == want  # the context still serves plain encodes in between
        oc.close()
    finally:
        ctx.close()


def test_multi_gpu_context(dictionary, oracle_ctx_factory):
    """mnenc_create_multi: ONE handle over several GPUs of this process (what the shim and the CLI use); plain
    mnenc_encode_rgb8 calls on it shard the image by block columns. Needs >= 2 GPUs."""
    import torch
    n = min(torch.cuda.device_count(), 4)
    if n < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mn.createCompressionContextFast(32, 8, 3.5, dictionary=dictionary, device=list(range(n)))
    try:
        oc = oracle_ctx_factory(ctx.Quant)
        for (w, h, seed) in ((512, 512, 1), (1023, 77, 2), (8, 8, 3)):
            img = synthetic_rgb(w, h, seed)
            assert mn.encodeImageFast(img, ctx) == oc.encode(img), (w, h)
        oc.close()
    finally:
        ctx.close()
