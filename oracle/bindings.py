"""TEST INFRASTRUCTURE — ctypes bindings for the two CPU checkers.

* ``RefLib``    -> oracle/_ref/libmnref.so : the UNMODIFIED reference (built by oracle/Makefile from /root/reference)
* ``OracleLib`` -> oracle/libmnoracle.so   : our CPU restatement (oracle/mn_oracle.cpp)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs may import this
module. The product package (imageexperiments_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libmnref.so")
ORACLE_SO = os.path.join(HERE, "libmnoracle.so")

_vp, _sz, _u64 = C.c_void_p, C.c_size_t, C.c_uint64


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def _take(lib_free, p, n, dtype) -> np.ndarray:
    """Copy n items out of a malloc'ed buffer and free it."""
    if n:
        out = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n * np.dtype(dtype).itemsize,)).copy().view(dtype)
    else:
        out = np.zeros(0, dtype)
    lib_free(p)
    return out


def ref_available() -> bool:
    return os.path.exists(REF_SO)


def oracle_available() -> bool:
    return os.path.exists(ORACLE_SO)


class Dictionary:
    """Dictionary floats exported from one reference context (import, never regenerate)."""

    def __init__(self, K, bs, base, detail_rows, detail, quant):
        self.K, self.bs = K, bs
        self.base = np.ascontiguousarray(base, np.float32)  # [n_base, bs*bs]
        self.detail_rows = np.ascontiguousarray(detail_rows, np.int32)  # [n_base]
        self.detail = [np.ascontiguousarray(d, np.float32) for d in detail]  # 3 x [sum rows, bs*bs]
        self.quant = np.ascontiguousarray(quant, np.float32)  # [3K]


class RefLib:
    def __init__(self, path: str = REF_SO):
        L = self.L = C.CDLL(path)
        L.mnref_last_error.restype = C.c_char_p
        L.mnref_create.restype = _vp
        L.mnref_create.argtypes = [C.c_int, C.c_int, C.c_double]
        L.mnref_destroy.argtypes = [_vp]
        L.mnref_free.argtypes = [_vp]
        L.mnref_quant_tables.argtypes = [C.c_int, C.c_int, C.c_double, _vp]
        L.mnref_get_quant.argtypes = [_vp, _vp]
        L.mnref_set_quant.argtypes = [_vp, _vp]
        L.mnref_dict_shape.argtypes = [_vp, _vp, _vp, _vp]
        L.mnref_get_base.argtypes = [_vp, _vp]
        L.mnref_get_detail.argtypes = [_vp, C.c_int, _vp]
        L.mnref_dynamic_probe.argtypes = [_vp, C.c_int, C.c_int, _vp, _vp]
        L.mnref_encode.argtypes = [_vp, _vp, _sz, _sz, _vp, _vp]
        L.mnref_symbols.argtypes = [_vp, _vp, _sz, _sz, _sz, _sz, _vp, _vp]
        L.mnref_parse.argtypes = [_vp, _sz] + [_vp] * 8
        L.mnref_decode.argtypes = [_vp, _sz, _vp, _vp, _vp]
        L.mnref_psnr.restype = C.c_double
        L.mnref_psnr.argtypes = [_vp, _vp, _sz, _sz]
        for f in ("mnref_write_huffman_or_golomb", "mnref_huffman_encode"):
            getattr(L, f).argtypes = [_vp, _sz, _vp, _vp, _vp]
        L.mnref_huffman_decode.argtypes = [_vp, _sz, _vp, _vp]
        L.mnref_rle_encode.argtypes = [_vp, _sz, _vp, _vp]
        L.mnref_rle_decode.argtypes = [_vp, _sz, _vp, _vp]
        L.mnref_golomb_length.restype = C.c_uint32
        L.mnref_golomb_length.argtypes = [C.c_uint32, C.c_uint32]
        L.mnref_golomb_write.argtypes = [_vp, _sz, C.c_uint32, _vp, _vp, _vp]
        L.mnref_elias_fano_length.restype = C.c_uint32
        L.mnref_elias_fano_length.argtypes = [_sz, C.c_uint16]
        L.mnref_elias_fano_write.argtypes = [_vp, _sz, C.c_uint16, _vp, _vp, _vp]
        L.mnref_zigzag_encode.restype = C.c_uint32
        L.mnref_zigzag_encode.argtypes = [C.c_int32]

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError("reference threw: " + self.L.mnref_last_error().decode())

    # -- context -------------------------------------------------------------------------------
    def create(self, K=32, bs=8, bpp=3.5):
        ctx = self.L.mnref_create(K, bs, float(bpp))
        if not ctx:
            raise RuntimeError("mnref_create failed: " + self.L.mnref_last_error().decode())
        return ctx

    def destroy(self, ctx):
        self.L.mnref_destroy(ctx)

    def quant_tables(self, K, bs, bpp) -> np.ndarray:
        q = np.zeros(3 * K, np.float32)
        self._check(self.L.mnref_quant_tables(K, bs, float(bpp), _ptr(q)))
        return q

    def get_quant(self, ctx, K=32) -> np.ndarray:
        q = np.zeros(3 * K, np.float32)
        self.L.mnref_get_quant(ctx, _ptr(q))
        return q

    def set_quant(self, ctx, q):
        q = np.ascontiguousarray(q, np.float32)
        self.L.mnref_set_quant(ctx, _ptr(q))

    def dictionary(self, ctx, K=32, bs=8) -> Dictionary:
        nb, cols = C.c_int(), C.c_int()
        self.L.mnref_dict_shape(ctx, C.byref(nb), C.byref(cols), None)
        rows = np.zeros(nb.value, np.int32)
        self.L.mnref_dict_shape(ctx, C.byref(nb), C.byref(cols), _ptr(rows))
        base = np.zeros((nb.value, cols.value), np.float32)
        self.L.mnref_get_base(ctx, _ptr(base))
        det = []
        for ch in range(3):
            d = np.zeros((int(rows.sum()), cols.value), np.float32)
            self.L.mnref_get_detail(ctx, ch, _ptr(d))
            det.append(d)
        return Dictionary(K, bs, base, rows, det, self.get_quant(ctx, K))

    def dynamic_probe(self, ctx, ch, prev, cols=64) -> np.ndarray:
        rows = C.c_int()
        self._check(self.L.mnref_dynamic_probe(ctx, ch, prev, None, C.byref(rows)))
        out = np.zeros((rows.value, cols), np.float32)
        self._check(self.L.mnref_dynamic_probe(ctx, ch, prev, _ptr(out), C.byref(rows)))
        return out

    # -- codec ---------------------------------------------------------------------------------
    def encode(self, ctx, rgb: np.ndarray) -> bytes:
        rgb = np.ascontiguousarray(rgb, np.uint8)
        h, w, _ = rgb.shape
        out, n = _vp(), _sz()
        self._check(self.L.mnref_encode(ctx, _ptr(rgb), w, h, C.byref(out), C.byref(n)))
        return _take(self.L.mnref_free, out, n.value, np.uint8).tobytes()

    def symbols(self, ctx, rgb: np.ndarray, K=32, bs=8, bx0=0, bx1=None) -> Tuple[np.ndarray, np.ndarray]:
        rgb = np.ascontiguousarray(rgb, np.uint8)
        h, w, _ = rgb.shape
        nbx, nby = (w + bs - 1) // bs, (h + bs - 1) // bs
        bx1 = nbx if bx1 is None else bx1
        nb = (bx1 - bx0) * nby
        counts = np.zeros((nb, 3), np.uint16)
        pairs = np.zeros((nb, 3, K, 2), np.uint16)
        self._check(self.L.mnref_symbols(ctx, _ptr(rgb), w, h, bx0, bx1, _ptr(counts), _ptr(pairs)))
        return counts, pairs

    def parse(self, data: bytes):
        buf = np.frombuffer(data, np.uint8)
        K, bs, w, h, nl = _sz(), _sz(), _sz(), _sz(), _sz()
        lengths, streams = _vp(), _vp()
        sizes = np.zeros(6 * 64, np.uint64)
        self._check(self.L.mnref_parse(_ptr(buf), len(data), C.byref(K), C.byref(bs), C.byref(w), C.byref(h), C.byref(lengths),
                                       C.byref(nl), _ptr(sizes), C.byref(streams)))
        sizes = sizes[: 6 * K.value].astype(np.int64)
        ln = _take(self.L.mnref_free, lengths, nl.value, np.uint16)
        flat = _take(self.L.mnref_free, streams, int(sizes.sum()), np.uint16)
        offs = np.concatenate([[0], np.cumsum(sizes)])
        return dict(K=K.value, bs=bs.value, w=w.value, h=h.value, lengths=ln,
                    streams=[flat[offs[i]:offs[i + 1]] for i in range(len(sizes))])

    def decode(self, data: bytes) -> np.ndarray:
        buf = np.frombuffer(data, np.uint8)
        out, w, h = _vp(), _sz(), _sz()
        self._check(self.L.mnref_decode(_ptr(buf), len(data), C.byref(out), C.byref(w), C.byref(h)))
        return _take(self.L.mnref_free, out, w.value * h.value * 3, np.uint8).reshape(h.value, w.value, 3)

    def psnr(self, a: np.ndarray, b: np.ndarray) -> float:
        a = np.ascontiguousarray(a, np.uint8)
        b = np.ascontiguousarray(b, np.uint8)
        h, w, _ = a.shape
        return float(self.L.mnref_psnr(_ptr(a), _ptr(b), w, h))

    # -- entropy primitives -----------------------------------------------------------------------
    def _bits_call(self, fn, data, *extra):
        out, nbits, nbytes = _vp(), _sz(), _sz()
        self._check(fn(_ptr(data), len(data), *extra, C.byref(out), C.byref(nbits), C.byref(nbytes)))
        return _take(self.L.mnref_free, out, nbytes.value, np.uint8).tobytes(), nbits.value

    def huffman_or_golomb(self, data) -> Tuple[bytes, int]:
        return self._bits_call(self.L.mnref_write_huffman_or_golomb, np.ascontiguousarray(data, np.uint16))

    def huffman(self, data) -> Tuple[bytes, int]:
        return self._bits_call(self.L.mnref_huffman_encode, np.ascontiguousarray(data, np.uint16))

    def huffman_decode(self, blob: bytes) -> np.ndarray:
        buf = np.frombuffer(blob, np.uint8)
        out, n = _vp(), _sz()
        self._check(self.L.mnref_huffman_decode(_ptr(buf), len(blob), C.byref(out), C.byref(n)))
        return _take(self.L.mnref_free, out, n.value, np.uint16)

    def rle(self, data) -> np.ndarray:
        data = np.ascontiguousarray(data, np.uint16)
        out, n = _vp(), _sz()
        self._check(self.L.mnref_rle_encode(_ptr(data), len(data), C.byref(out), C.byref(n)))
        return _take(self.L.mnref_free, out, n.value, np.uint16)

    def rle_decode(self, data) -> np.ndarray:
        data = np.ascontiguousarray(data, np.uint16)
        out, n = _vp(), _sz()
        self._check(self.L.mnref_rle_decode(_ptr(data), len(data), C.byref(out), C.byref(n)))
        return _take(self.L.mnref_free, out, n.value, np.uint16)

    def golomb_length(self, v, m) -> int:
        return int(self.L.mnref_golomb_length(int(v), int(m)))

    def golomb(self, values, m) -> Tuple[bytes, int]:
        return self._bits_call(self.L.mnref_golomb_write, np.ascontiguousarray(values, np.uint32), C.c_uint32(m))

    def elias_fano_bound(self, n, max_symbol) -> int:
        return int(self.L.mnref_elias_fano_length(n, max_symbol))

    def elias_fano(self, seq, max_symbol) -> Tuple[bytes, int]:
        return self._bits_call(self.L.mnref_elias_fano_write, np.ascontiguousarray(seq, np.uint16), C.c_uint16(max_symbol))

    def zigzag(self, x) -> int:
        return int(self.L.mnref_zigzag_encode(int(x)))


class OracleLib:
    """Our CPU restatement (oracle/mn_oracle.cpp). Needs a Dictionary exported from a reference build."""

    def __init__(self, path: str = ORACLE_SO):
        L = self.L = C.CDLL(path)
        L.mno_dict_create.restype = _vp
        L.mno_dict_create.argtypes = [C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp]
        L.mno_dict_set_quant.argtypes = [_vp, _vp]
        L.mno_dict_destroy.argtypes = [_vp]
        L.mno_free.argtypes = [_vp]
        L.mno_gather_block.argtypes = [_vp, _sz, _sz, _sz, _sz, _sz, C.c_int, _vp]
        L.mno_mp_block.restype = C.c_int
        L.mno_mp_block.argtypes = [_vp, C.c_int, _vp, _vp]
        L.mno_symbols.argtypes = [_vp, _vp, _sz, _sz, _sz, _sz, _sz, _vp, _vp]
        L.mno_rle.restype = _vp
        L.mno_rle.argtypes = [_vp, _sz, _vp]
        for f in ("mno_huffman", "mno_huffman_or_golomb"):
            getattr(L, f).restype = _vp
            getattr(L, f).argtypes = [_vp, _sz, _vp, _vp]
        L.mno_golomb_length.restype = C.c_uint32
        L.mno_golomb_length.argtypes = [C.c_uint32, C.c_uint32]
        L.mno_golomb.restype = _vp
        L.mno_golomb.argtypes = [_vp, _sz, C.c_uint32, _vp, _vp]
        L.mno_elias_fano_bound.restype = C.c_uint32
        L.mno_elias_fano_bound.argtypes = [_sz, C.c_uint16]
        L.mno_elias_fano.restype = _vp
        L.mno_elias_fano.argtypes = [_vp, _sz, C.c_uint16, _vp, _vp]
        L.mno_zigzag.restype = C.c_uint32
        L.mno_zigzag.argtypes = [C.c_int32]
        L.mno_write_file.restype = _vp
        L.mno_write_file.argtypes = [C.c_int, C.c_int, _sz, _sz, _vp, _vp, _vp, _vp]
        L.mno_encode.restype = _vp
        L.mno_encode.argtypes = [_vp, _vp, _sz, _sz, _sz, _vp]
        L.mno_parse.restype = C.c_int
        L.mno_parse.argtypes = [_vp, _sz] + [_vp] * 8
        L.mno_decode.restype = C.c_int
        L.mno_decode.argtypes = [_vp, _vp, _sz, _vp, _vp, _vp]

    def parse(self, data: bytes):
        """Same result layout as RefLib.parse; raises ValueError on an invalid stream."""
        buf = np.frombuffer(data, np.uint8)
        K, bs, w, h, nl = _sz(), _sz(), _sz(), _sz(), _sz()
        lengths, streams = _vp(), _vp()
        sizes = np.zeros(6 * 64, np.uint64)
        if self.L.mno_parse(_ptr(buf), len(data), C.byref(K), C.byref(bs), C.byref(w), C.byref(h), C.byref(lengths), C.byref(nl),
                            _ptr(sizes), C.byref(streams)):
            raise ValueError("invalid .mn stream")
        sizes = sizes[: 6 * K.value].astype(np.int64)
        ln = _take(self.L.mno_free, lengths, nl.value, np.uint16)
        flat = _take(self.L.mno_free, streams, int(sizes.sum()), np.uint16)
        offs = np.concatenate([[0], np.cumsum(sizes)])
        return dict(K=K.value, bs=bs.value, w=w.value, h=h.value, lengths=ln,
                    streams=[flat[offs[i]:offs[i + 1]] for i in range(len(sizes))])

    # -- dictionary handle ---------------------------------------------------------------------------
    def create(self, d: Dictionary):
        h = self.L.mno_dict_create(d.K, d.bs, d.base.shape[0], _ptr(d.base), _ptr(d.detail_rows), _ptr(d.detail[0]),
                                   _ptr(d.detail[1]), _ptr(d.detail[2]), _ptr(d.quant))
        if not h:
            raise RuntimeError("mno_dict_create failed")
        return _OracleCtx(self, h, d)

    # -- entropy primitives -----------------------------------------------------------------------------
    def _bits(self, fn, data, *extra):
        nbits, nbytes = _u64(), _u64()
        p = fn(_ptr(data), len(data), *extra, C.byref(nbits), C.byref(nbytes))
        return _take(self.L.mno_free, p, nbytes.value, np.uint8).tobytes(), nbits.value

    def rle(self, data) -> np.ndarray:
        data = np.ascontiguousarray(data, np.uint16)
        n = _sz()
        p = self.L.mno_rle(_ptr(data), len(data), C.byref(n))
        return _take(self.L.mno_free, p, n.value, np.uint16)

    def huffman(self, data):
        return self._bits(self.L.mno_huffman, np.ascontiguousarray(data, np.uint16))

    def huffman_or_golomb(self, data):
        return self._bits(self.L.mno_huffman_or_golomb, np.ascontiguousarray(data, np.uint16))

    def golomb_length(self, v, m) -> int:
        return int(self.L.mno_golomb_length(int(v), int(m)))

    def golomb(self, values, m):
        return self._bits(self.L.mno_golomb, np.ascontiguousarray(values, np.uint32), C.c_uint32(m))

    def elias_fano_bound(self, n, max_symbol) -> int:
        return int(self.L.mno_elias_fano_bound(n, max_symbol))

    def elias_fano(self, seq, max_symbol):
        return self._bits(self.L.mno_elias_fano, np.ascontiguousarray(seq, np.uint16), C.c_uint16(max_symbol))

    def zigzag(self, x) -> int:
        return int(self.L.mno_zigzag(int(x)))

    def write_file(self, K, bs, w, h, quant, counts, pairs) -> bytes:
        quant = np.ascontiguousarray(quant, np.float32)
        counts = np.ascontiguousarray(counts, np.uint16)
        pairs = np.ascontiguousarray(pairs, np.uint16)
        n = _u64()
        p = self.L.mno_write_file(K, bs, w, h, _ptr(quant), _ptr(counts), _ptr(pairs), C.byref(n))
        return _take(self.L.mno_free, p, n.value, np.uint8).tobytes()


class _OracleCtx:
    def __init__(self, lib: OracleLib, h, d: Dictionary):
        self.lib, self.h, self.d = lib, h, d  # keep d alive: the C side borrows its arrays

    def close(self):
        if self.h:
            self.lib.L.mno_dict_destroy(self.h)
            self.h = None

    def set_quant(self, q):
        q = np.ascontiguousarray(q, np.float32)
        self.d.quant = q
        self.lib.L.mno_dict_set_quant(self.h, _ptr(q))

    def gather_block(self, rgb, x0, y0) -> np.ndarray:
        rgb = np.ascontiguousarray(rgb, np.uint8)
        h, w, _ = rgb.shape
        out = np.zeros((3, self.d.bs * self.d.bs), np.float32)
        self.lib.L.mno_gather_block(_ptr(rgb), w, h, 3 * w, x0, y0, self.d.bs, _ptr(out))
        return out

    def mp_block(self, ch, vec):
        vec = np.ascontiguousarray(vec, np.float32)
        pairs = np.zeros((self.d.K, 2), np.uint16)
        count = self.lib.L.mno_mp_block(self.h, ch, _ptr(vec), _ptr(pairs))
        return count, pairs

    def symbols(self, rgb, bx0=0, bx1=None):
        rgb = np.ascontiguousarray(rgb, np.uint8)
        h, w, _ = rgb.shape
        bs, K = self.d.bs, self.d.K
        nbx, nby = (w + bs - 1) // bs, (h + bs - 1) // bs
        bx1 = nbx if bx1 is None else bx1
        nb = (bx1 - bx0) * nby
        counts = np.zeros((nb, 3), np.uint16)
        pairs = np.zeros((nb, 3, K, 2), np.uint16)
        self.lib.L.mno_symbols(self.h, _ptr(rgb), w, h, 3 * w, bx0, bx1, _ptr(counts), _ptr(pairs))
        return counts, pairs

    def symbols_parallel(self, rgb, workers: Optional[int] = None):
        """symbols() of the whole image, block columns spread over host threads (mno_symbols is re-entrant and ctypes
        releases the GIL): the full 4928x3264 image takes seconds on the GPU box's cores instead of ~100 s."""
        from concurrent.futures import ThreadPoolExecutor
        rgb = np.ascontiguousarray(rgb, np.uint8)
        h, w, _ = rgb.shape
        bs = self.d.bs
        nbx = (w + bs - 1) // bs
        workers = workers or max(1, len(os.sched_getaffinity(0)))
        step = max(1, -(-nbx // (4 * workers)))
        ranges = [(b, min(b + step, nbx)) for b in range(0, nbx, step)]
        with ThreadPoolExecutor(workers) as ex:
            parts = list(ex.map(lambda r: self.symbols(rgb, r[0], r[1]), ranges))
        return np.concatenate([p[0] for p in parts]), np.concatenate([p[1] for p in parts])

    def encode(self, rgb) -> bytes:
        rgb = np.ascontiguousarray(rgb, np.uint8)
        h, w, _ = rgb.shape
        n = _u64()
        p = self.lib.L.mno_encode(self.h, _ptr(rgb), w, h, 3 * w, C.byref(n))
        return _take(self.lib.L.mno_free, p, n.value, np.uint8).tobytes()

    def decode(self, data: bytes) -> np.ndarray:
        buf = np.frombuffer(data, np.uint8)
        out, w, h = _vp(), _sz(), _sz()
        rc = self.lib.L.mno_decode(self.h, _ptr(buf), len(data), C.byref(out), C.byref(w), C.byref(h))
        if rc:
            raise ValueError("invalid .mn stream" if rc == -1 else "block size differs from the dictionary's")
        return _take(self.lib.L.mno_free, out, w.value * h.value * 3, np.uint8).reshape(h.value, w.value, 3)
