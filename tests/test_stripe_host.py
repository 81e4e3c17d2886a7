"""Multi-GPU stripe path, host logic, on the CPU with world_size 2 and 3 (gloo): every rank holds a contiguous slice of
each symbol stream, all-gathers boundary records, stitches runs with the product's host code (mnenc_debug_stitch_runs)
and run-length codes its slice alone; the concatenation over ranks must equal the oracle's RLE of the whole stream
(huffman::runLengthEncode, Huffman.cpp:246-279), including runs that cross ranks and the 32 769 chunk rule."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHUNK = 32769


def make_streams():
    rng = np.random.default_rng(42)
    streams = [
        np.zeros(0, np.uint16),
        np.full(7, 5, np.uint16),
        np.full(100000, 9, np.uint16),                                  # one run over all ranks, three chunks
        np.concatenate([rng.integers(0, 3, 50), np.full(70000, 2), rng.integers(0, 3, 50)]).astype(np.uint16),
        rng.integers(0, 2, 4000).astype(np.uint16),                     # many short runs
        np.repeat(rng.integers(0, 50, 300), rng.integers(1, 400, 300)).astype(np.uint16),
        np.concatenate([np.full(32769, 1), np.full(32770, 1)]).astype(np.uint16),
        np.array([0, 0, 0, 1], np.uint16),                              # leading zeros are a normal run
    ]
    return streams


def cuts(n, world, rng):
    """Slice boundaries, deliberately ragged (empty slices, cuts inside runs)."""
    c = np.sort(rng.integers(0, n + 1, world - 1)) if n else np.zeros(world - 1, int)
    return [0] + [int(x) for x in c] + [n]


def rle_size(n):
    """RLE symbols of one run of n elements that lies wholly inside a slice."""
    full, rem = divmod(n, CHUNK)
    return 3 * full + (0 if rem == 0 else 1 if rem == 1 else 3)


def boundary(v):
    """The record a rank publishes per stream slice (mnenc_stripe_rle's rec2 row)."""
    if v.size == 0:
        return [0, 0, 0, 0, 0, 0]
    ch = np.nonzero(v[1:] != v[:-1])[0]
    first_run = int(ch[0] + 1) if ch.size else v.size
    last_run = int(v.size - (ch[-1] + 1)) if ch.size else v.size
    starts = np.concatenate([[0], ch + 1, [v.size]])
    interior = sum(rle_size(int(starts[i + 1] - starts[i])) for i in range(1, len(starts) - 2))
    return [v.size, int(v[0]), first_run, int(v[-1]), last_run, interior]


def rle_slice(v, carry, ext):
    """What the stripe RLE kernels emit for one slice: a chunk [v, v, len-2] / [v] belongs to the rank holding its first
    element; chunks are counted from the run's global start."""
    out = []
    if v.size == 0:
        return out
    starts = np.concatenate([[0], np.nonzero(v[1:] != v[:-1])[0] + 1, [v.size]])
    for i in range(len(starts) - 1):
        s0, s1 = int(starts[i]), int(starts[i + 1])
        c = carry if s0 == 0 else 0
        total = c + (s1 - s0) + (ext if s1 == v.size else 0)
        k = (c + CHUNK - 1) // CHUNK
        while k * CHUNK < c + (s1 - s0):
            ln = min(CHUNK, total - k * CHUNK)
            out += [int(v[s0])] if ln == 1 else [int(v[s0]), int(v[s0]), ln - 2]
            k += 1
    return out


def worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from imageexperiments_b200 import _lib
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    L = _lib.load()
    streams = make_streams()
    rng = np.random.default_rng(7)
    ns = len(streams)
    slices = []
    for v in streams:
        c = cuts(v.size, world, rng)
        slices.append(v[c[rank]:c[rank + 1]])
    rec = torch.tensor([boundary(v) for v in slices], dtype=torch.int64)
    allrec = [torch.zeros_like(rec) for _ in range(world)]
    dist.all_gather(allrec, rec)
    records = np.ascontiguousarray(torch.stack(allrec).numpy().astype(np.uint32))
    carry = np.zeros(ns, np.uint32)
    ext = np.zeros(ns, np.uint32)
    rc = L.mnenc_debug_stitch_runs(records.ctypes.data, world, ns, rank, carry.ctypes.data, ext.ctypes.data)
    assert rc == 0
    mine = [rle_slice(v, int(carry[i]), int(ext[i])) for i, v in enumerate(slices)]
    # the size the device scan would use must agree with what is emitted
    for i, v in enumerate(slices):
        if v.size and (v == v[0]).all():
            want = L.mnenc_debug_rle_chunk_size(int(carry[i]), int(v.size), int(carry[i]) + int(v.size) + int(ext[i]))
            assert want == len(mine[i]), (i, want, len(mine[i]))
    # what every rank derives from the records alone: per-rank emitted sizes, whole-stream sizes, RLE-or-raw decision
    emitted = np.zeros((world, ns), np.uint32)
    use_rle = np.zeros(ns, np.uint8)
    raw_len = np.zeros(ns, np.uint64)
    rle_len = np.zeros(ns, np.uint64)
    assert L.mnenc_debug_plan_stripes(records.ctypes.data, world, ns, emitted.ctypes.data, use_rle.ctypes.data, raw_len.ctypes.data,
                                      rle_len.ctypes.data) == 0
    assert [int(x) for x in emitted[rank]] == [len(m) for m in mine]
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    if rank == 0:
        q.put(([sum((g[i] for g in gathered), []) for i in range(ns)], use_rle.tolist(), raw_len.tolist(), rle_len.tolist()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_rle_stitching_matches_oracle(world, oraclelib):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got, use_rle, raw_len, rle_len = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for i, (v, g) in enumerate(zip(make_streams(), got)):
        want = oraclelib.rle(v)
        assert np.array_equal(np.asarray(g, np.uint16), want), (v.size, len(g), want.size)
        assert raw_len[i] == v.size and rle_len[i] == want.size
        assert bool(use_rle[i]) == (want.size + 4 < v.size)   # CompressedImage.cpp:508-516


@pytest.mark.parametrize("nranks", [1, 2, 3, 8])
def test_sharded_code_tables_equal_unsharded(nranks):
    """In-library stripe exchange (csrc/stripe_comm.cu planShardBuild / planShardApply): the table builds dealt over the ranks,
    serialised and applied on a rank that holds its entries in another order must give the codes, lengths, Golomb parameters
    and payload sizes of the plain per-sequence build (mnenc_debug_plan_sequence)."""
    import ctypes as C

    from imageexperiments_b200 import _lib
    from imageexperiments_b200.compressed import planSequenceHost
    rng = np.random.default_rng(1234 + nranks)
    n_seq = 25
    syms, freqs, first, start = [], [], [], [0]
    for q in range(n_seq):
        n = [0, 1, 2, 7, 300, 2000][q % 6] if q < 18 else int(rng.integers(1, 400))
        s = rng.choice(65535, size=n, replace=False).astype(np.uint32)
        # geometric-like counts for some sequences (Golomb wins), flat ones for others (Huffman wins)
        f = (rng.geometric(0.02, size=n) if q % 2 else rng.integers(1, 50, size=n)).astype(np.uint32)
        syms.append(s), freqs.append(f), first.append(rng.permutation(n).astype(np.uint32))
        start.append(start[-1] + n)
    symbols, fr, fp = np.concatenate(syms), np.concatenate(freqs), np.concatenate(first)
    start = np.asarray(start, np.uint32)
    n = int(start[-1])
    codes, lens = np.zeros(n, np.uint32), np.zeros(n, np.uint8)
    bits, gm = np.zeros(n_seq, np.uint64), np.zeros(n_seq, np.uint32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    _lib.check(_lib.load().mnenc_debug_plan_sharded(p(symbols), p(fr), p(fp), p(start), n_seq, nranks, p(codes), p(lens), p(bits), p(gm)))
    for q in range(n_seq):
        want = planSequenceHost(syms[q], freqs[q], first[q])
        a, b = int(start[q]), int(start[q + 1])
        assert int(gm[q]) == want["golomb_m"], q
        assert int(bits[q]) == want["payload_bits"], q
        if want["golomb_m"] == 0:
            assert np.array_equal(codes[a:b], want["codes"]) and np.array_equal(lens[a:b], want["lens"]), q
