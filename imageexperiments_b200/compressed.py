"""Host-side mirror of the reference's encoder interface (CompressionLib/inc/CompressedImage.h:54-76), backed by
the sm_100a library through the C ABI of include/mnenc.h.

    ctx = createCompressionContextFast(32, 8, 3.5)         # CompressedImage.cpp:317
    data = encodeImageFast(rgb, ctx)                       # CompressedImage.cpp:578 -> bytes of the .mn stream

Names and argument meaning follow the reference; Python conveniences: images are HxWx3 uint8 numpy arrays (or
CUDA torch tensors for the device entry point), the quality parameter is ``bppAllocation`` (a float, or "max"
as in Compression.cpp:104-114), and errors are exceptions (MnencError) instead of thrown pointers.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple, Union

import numpy as np

from . import _lib
from .dictionary import Dictionary

_vp = C.c_void_p


def _p(a: np.ndarray):
    return a.ctypes.data_as(_vp)


def createQuantizationTables(K: int, blockSize: int, bppAllocation: float) -> np.ndarray:
    """float overload of compressed::createQuantizationTables (CompressedImage.cpp:168-210) -> [3K] (Y,U,V)."""
    q = np.zeros(3 * K, np.float32)
    _lib.check(_lib.load().mnenc_quant_tables(K, blockSize, float(bppAllocation), _p(q)))
    return q


class CompressionContextFast:
    """compressed::CompressionContextFast (CompressedImage.h:46-55) living on one B200."""

    def __init__(self, K: int, blockSize: int, quant: np.ndarray, dictionary: Dictionary, device=0):
        """``device``: a CUDA ordinal, or a list of ordinals for ONE context over several GPUs of this process
        (mnenc_create_multi): encodeImageFast on it shards every image by block columns over those GPUs."""
        self.devices = [int(x) for x in device] if isinstance(device, (list, tuple)) else None
        self.K, self.BlockSize, self.device = K, blockSize, (self.devices[0] if self.devices else int(device))
        self.dictionary = dictionary
        self._h = _vp()
        L = _lib.load()
        d = dictionary
        if self.devices:
            devs = np.asarray(self.devices, np.int32)
            _lib.check(L.mnenc_create_multi(C.byref(self._h), _p(devs), len(self.devices), K, blockSize, _p(d.base), d.n_base, _p(d.detail_rows),
                                            _p(d.detail[0]), _p(d.detail[1]), _p(d.detail[2])))
        else:
            _lib.check(L.mnenc_create(C.byref(self._h), self.device, K, blockSize, _p(d.base), d.n_base, _p(d.detail_rows), _p(d.detail[0]),
                                      _p(d.detail[1]), _p(d.detail[2])))
        self.setQuant(quant)

    def setQuant(self, quant) -> None:
        self.Quant = np.ascontiguousarray(quant, np.float32).reshape(3 * self.K).copy()
        _lib.check(_lib.load().mnenc_set_quant(self._h, _p(self.Quant)))

    def close(self) -> None:
        if self._h:
            _lib.load().mnenc_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- accounting of the last encode --------------------------------------------------------------
    def last_timing(self) -> Tuple[np.ndarray, int]:
        ms = np.zeros(8, np.float32)
        n = _lib.check(_lib.load().mnenc_last_timing(self._h, _p(ms)))
        return ms, n

    def last_plan(self) -> Tuple[int, int, int]:
        """(sequences written with a Huffman table, as Golomb codes, symbol streams run-length coded) of the last encode."""
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        _lib.check(_lib.load().mnenc_last_plan(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def last_work(self) -> Tuple[float, int, int]:
        f, s, b = C.c_double(), C.c_uint64(), C.c_uint64()
        _lib.check(_lib.load().mnenc_last_work(self._h, C.byref(f), C.byref(s), C.byref(b)))
        return f.value, s.value, b.value


def createCompressionContextFast(K: int = 32, blockSize: int = 8, bppAllocation: Union[float, str] = 3.5,
                                 dictionary: Optional[Dictionary] = None, device=0) -> CompressionContextFast:
    """compressed::createCompressionContextFast (CompressedImage.cpp:317-341); the dictionary is imported data."""
    if dictionary is None:
        dictionary = Dictionary.load()
    elif isinstance(dictionary, str) and dictionary == "builtin":  # computed by the library: its own codec revision (Dictionary.build)
        dictionary = Dictionary.build(blockSize)
    if isinstance(bppAllocation, str):
        if bppAllocation != "max":
            raise ValueError("quality is a number or 'max'")
        quant = np.ones(3 * K, np.float32)  # Compression.cpp:104-114: "max" sets every quantiser to 1
    else:
        quant = createQuantizationTables(K, blockSize, bppAllocation)
    return CompressionContextFast(K, blockSize, quant, dictionary, device)


class MnBuffer:
    """A `.mn` stream still owned by the library (malloc'ed by the encoder, released on garbage collection): zero-copy
    alternative to ``bytes`` for streams of tens of MB. ``memoryview(buf.array)`` / ``bytes(buf)`` / ``len(buf)``."""

    def __init__(self, ptr: _vp, n: int):
        self._ptr, self._n = ptr, n
        self.array = (C.c_uint8 * n).from_address(ptr.value) if n else (C.c_uint8 * 0)()

    def __len__(self):
        return self._n

    def __bytes__(self):
        return C.string_at(self._ptr, self._n)

    def __eq__(self, other):
        return bytes(self) == bytes(other)

    def __del__(self):
        try:
            if self._ptr:
                _lib.load().mnenc_free(self._ptr)
                self._ptr = None
        except Exception:
            pass


def _take(out: _vp, n: int) -> bytes:
    try:
        return C.string_at(out, n)
    finally:
        _lib.load().mnenc_free(out)


def encodeImageFast(imgIn, context: CompressionContextFast, zero_copy: bool = False):
    """compressed::encodeImageFast (CompressedImage.cpp:578-633). ``imgIn``: HxWx3 uint8 numpy array on the host
    (copied to the GPU inside the call) or a CUDA torch uint8 tensor of the same shape (already resident). Returns the
    `.mn` stream as bytes (``zero_copy``: as an MnBuffer over the library's pinned host buffer)."""
    L = _lib.load()
    out, n = _vp(), C.c_size_t()
    if isinstance(imgIn, np.ndarray):
        if imgIn.dtype != np.uint8 or imgIn.ndim != 3 or imgIn.shape[2] != 3:
            raise ValueError("expected HxWx3 uint8")
        if not imgIn.flags.c_contiguous:
            imgIn = np.ascontiguousarray(imgIn)
        h, w, _ = imgIn.shape
        _lib.check(L.mnenc_encode_rgb8(context._h, _p(imgIn), w, h, imgIn.strides[0], C.byref(out), C.byref(n)))
    else:  # torch tensor
        if not imgIn.is_cuda:
            if str(imgIn.dtype) != "torch.uint8" or imgIn.dim() != 3 or imgIn.shape[2] != 3 or not imgIn.is_contiguous():
                raise ValueError("expected contiguous HxWx3 uint8")
            h, w, _ = imgIn.shape
            _lib.check(L.mnenc_encode_rgb8(context._h, _vp(imgIn.data_ptr()), w, h, 3 * w, C.byref(out), C.byref(n)))
        else:
            if str(imgIn.dtype) != "torch.uint8" or imgIn.dim() != 3 or imgIn.shape[2] != 3 or not imgIn.is_contiguous():
                raise ValueError("expected contiguous HxWx3 uint8")
            if imgIn.device.index != context.device:
                raise ValueError("tensor is on another device than the context")
            h, w, _ = imgIn.shape
            _lib.check(L.mnenc_encode_rgb8_device(context._h, _vp(imgIn.data_ptr()), w, h, 3 * w, C.byref(out), C.byref(n)))
    return MnBuffer(out, n.value) if zero_copy else _take(out, n.value)


def encodeImageBatch(images, context: CompressionContextFast, zero_copy: bool = False) -> list:
    """Throughput mode: encodeImageFast of every image of ``images`` (all host numpy arrays / CPU tensors, or all CUDA
    tensors on the context's device), pipelined over the context's lanes (mnenc_encode_batch_rgb8). Returns a list of
    bytes, each identical to encodeImageFast(image, context) — or, with ``zero_copy``, of MnBuffer objects over the
    library's own (pinned host) result buffers: copying 16 x 3.4 MB into Python bytes costs ~8 ms with the GPU idle."""
    n = len(images)
    if n == 0:
        return []
    ptrs, ws, hs, strides = (_vp * n)(), (C.c_size_t * n)(), (C.c_size_t * n)(), (C.c_size_t * n)()
    keep = []
    on_device = None
    for i, im in enumerate(images):
        if isinstance(im, np.ndarray):
            if im.dtype != np.uint8 or im.ndim != 3 or im.shape[2] != 3:
                raise ValueError("expected HxWx3 uint8")
            if not im.flags.c_contiguous:
                im = np.ascontiguousarray(im)
            dev, ptr = False, im.ctypes.data
        else:
            if str(im.dtype) != "torch.uint8" or im.dim() != 3 or im.shape[2] != 3 or not im.is_contiguous():
                raise ValueError("expected contiguous HxWx3 uint8")
            dev, ptr = bool(im.is_cuda), im.data_ptr()
            if dev and im.device.index != context.device:
                raise ValueError("tensor is on another device than the context")
        if on_device is None:
            on_device = dev
        elif on_device != dev:
            raise ValueError("a batch is either all host or all device images")
        keep.append(im)
        ptrs[i], hs[i], ws[i], strides[i] = ptr, im.shape[0], im.shape[1], 3 * im.shape[1]
    outs, sizes = (_vp * n)(), (C.c_size_t * n)()
    _lib.check(_lib.load().mnenc_encode_batch_rgb8(context._h, n, ptrs, int(on_device), ws, hs, strides, outs, sizes))
    if zero_copy:
        return [MnBuffer(_vp(outs[i]), sizes[i]) for i in range(n)]
    return [_take(_vp(outs[i]), sizes[i]) for i in range(n)]


class _PinnedBlock:
    """Owner of one library result buffer; released (recycled by the library) when the last view of it goes away."""

    def __init__(self, ptr: _vp):
        self._ptr = ptr

    def __del__(self):
        try:
            if self._ptr:
                _lib.load().mnenc_free(self._ptr)
                self._ptr = None
        except Exception:
            pass


def _pixels(rgb: _vp, w: int, h: int, zero_copy: bool) -> np.ndarray:
    n = 3 * w * h
    arr = (C.c_uint8 * n).from_address(rgb.value)
    arr._owner = _PinnedBlock(rgb)  # numpy keeps the ctypes array alive, the array keeps the pinned block from being recycled
    img = np.frombuffer(arr, np.uint8).reshape(h, w, 3)
    return img if zero_copy else img.copy()


def calculatePSNR(original: np.ndarray, decoded: np.ndarray) -> float:
    """compressed::calculatePSNR (CompressedImage.cpp:343-357): 20 log10(3 * 255) - 10 log10(sum over R,G,B of squared error
    / (width * height)) — the reference's own (non-standard) peak, kept so that figures compare with its `-n` output. Host."""
    a = np.asarray(original, np.float64)
    b = np.asarray(decoded, np.float64)
    if a.shape != b.shape or a.ndim != 3 or a.shape[2] != 3:
        raise ValueError("expected two HxWx3 images of the same size")
    mse = float(((a - b) ** 2).sum()) / (a.shape[0] * a.shape[1])
    return float(20.0 * np.log10(3.0 * 255.0) - 10.0 * np.log10(mse)) if mse > 0 else float("inf")


def _stream(data):
    """bytes / bytearray / MnBuffer / uint8 array -> (pointer, size, keep-alive)."""
    if isinstance(data, MnBuffer):
        return C.cast(data.array, _vp), len(data), data
    a = np.frombuffer(data, np.uint8) if not isinstance(data, np.ndarray) else np.ascontiguousarray(data, np.uint8)
    return _p(a), a.size, a


def decodeImageFast(data, context: CompressionContextFast, out=None, zero_copy: bool = False):
    """compressed::decodeImageFast (CompressedImage.cpp:837-889): `.mn` stream -> HxWx3 uint8 pixels, bit for bit those of the
    reference's decoder. ``data``: bytes, MnBuffer or uint8 array. The dictionary is the context's, the quantisation tables the
    stream's. ``out``: optional CUDA torch uint8 tensor [H, W, 3] (contiguous) that receives the pixels on the GPU (returned);
    otherwise a numpy array (``zero_copy``: a view of the library's pinned buffer, kept alive by the array)."""
    L = _lib.load()
    ptr, n, keep = _stream(data)
    w, h = C.c_size_t(), C.c_size_t()
    if out is not None:
        if not out.is_cuda or str(out.dtype) != "torch.uint8" or out.dim() != 3 or out.shape[2] != 3 or not out.is_contiguous():
            raise ValueError("expected a contiguous CUDA HxWx3 uint8 tensor")
        if out.device.index != context.device:
            raise ValueError("tensor is on another device than the context")
        _lib.check(L.mnenc_decode_rgb8_device(context._h, ptr, n, _vp(out.data_ptr()), out.numel(), 3 * out.shape[1], C.byref(w), C.byref(h)))
        if (h.value, w.value) != tuple(out.shape[:2]):
            raise ValueError(f"stream holds a {w.value}x{h.value} image, the tensor is {out.shape[1]}x{out.shape[0]}")
        return out
    rgb = _vp()
    _lib.check(L.mnenc_decode_rgb8(context._h, ptr, n, C.byref(rgb), C.byref(w), C.byref(h)))
    return _pixels(rgb, w.value, h.value, zero_copy)


def decodeImageBatch(streams, context: CompressionContextFast, zero_copy: bool = False) -> list:
    """Throughput mode: decodeImageFast of every stream of ``streams`` (bytes / MnBuffer / uint8 arrays), pipelined over the
    context's lanes (mnenc_decode_batch_rgb8). Returns a list of HxWx3 uint8 arrays, each identical to
    decodeImageFast(stream, context)."""
    n = len(streams)
    if n == 0:
        return []
    ptrs, sizes = (_vp * n)(), (C.c_size_t * n)()
    keep = []
    for i, d in enumerate(streams):
        ptr, size, k = _stream(d)
        keep.append(k)
        ptrs[i], sizes[i] = C.cast(ptr, _vp).value, size
    outs, ws, hs = (_vp * n)(), (C.c_size_t * n)(), (C.c_size_t * n)()
    _lib.check(_lib.load().mnenc_decode_batch_rgb8(context._h, n, ptrs, sizes, outs, ws, hs))
    return [_pixels(_vp(outs[i]), ws[i], hs[i], zero_copy) for i in range(n)]


def streamInfo(data) -> dict:
    """Header of a `.mn` stream (CompressedImage.cpp:721-736): width, height, K, blockSize, Quant [3K]. Host only."""
    ptr, n, keep = _stream(data)
    w, h, K, bs = C.c_size_t(), C.c_size_t(), C.c_int(), C.c_int()
    q = np.zeros(96, np.float32)
    _lib.check(_lib.load().mnenc_stream_info(ptr, n, C.byref(w), C.byref(h), C.byref(K), C.byref(bs), _p(q)))
    return dict(width=w.value, height=h.value, K=K.value, blockSize=bs.value, Quant=q[: 3 * K.value].copy())


def parseStream(data, threads: int = 1) -> dict:
    """Debug tap, host only: compressed::readCompressedFast (CompressedImage.cpp:709-782) -> lengths [3*nblocks] and the 6K
    symbol streams (run lengths expanded, DC chains undone). ``threads`` > 1 uses the decoder's parallel reader (same result)."""
    L = _lib.load()
    info = streamInfo(data)
    ptr, n, keep = _stream(data)
    nb = ((info["width"] + 7) // 8) * ((info["height"] + 7) // 8)
    lengths = np.zeros(3 * nb, np.uint16)
    offs = np.zeros(6 * info["K"] + 1, np.uint64)
    _lib.check(L.mnenc_debug_parse(ptr, n, _p(lengths), lengths.size, _p(offs), None, 0, threads))
    symbols = np.zeros(int(offs[-1]) + 1, np.uint16)
    _lib.check(L.mnenc_debug_parse(ptr, n, _p(lengths), lengths.size, _p(offs), _p(symbols), symbols.size, threads))
    o = offs.astype(np.int64)
    return dict(info, lengths=lengths, streams=[symbols[o[i]:o[i + 1]] for i in range(len(o) - 1)])


def lastSymbols(context: CompressionContextFast, width: int, height: int) -> Tuple[np.ndarray, np.ndarray]:
    """Debug tap: (counts [nblocks,3], pairs [nblocks,3,K,2]) of the last encode, column-major block order."""
    nb = ((width + 7) // 8) * ((height + 7) // 8)
    counts = np.zeros((nb, 3), np.uint16)
    pairs = np.zeros((nb, 3, context.K, 2), np.uint16)
    _lib.check(_lib.load().mnenc_get_symbols(context._h, _p(counts), _p(pairs)))
    return counts, pairs


def packSymbols(context: CompressionContextFast, width: int, height: int, counts: np.ndarray, pairs: np.ndarray) -> bytes:
    """Back end only (compressed::writeCompressed, CompressedImage.cpp:462-519) on given symbols."""
    counts = np.ascontiguousarray(counts, np.uint16)
    pairs = np.ascontiguousarray(pairs, np.uint16)
    nb = ((width + 7) // 8) * ((height + 7) // 8)
    if counts.size != 3 * nb or pairs.size != 3 * nb * context.K * 2:
        raise ValueError("symbol arrays do not match the image geometry")
    out, n = _vp(), C.c_size_t()
    _lib.check(_lib.load().mnenc_pack_symbols(context._h, width, height, _p(counts), _p(pairs), C.byref(out), C.byref(n)))
    return _take(out, n.value)


def planSequenceHost(symbols, freqs, first_pos):
    """Host table builder test hook (mnenc_debug_plan_sequence): returns a dict describing the coding plan."""
    symbols = np.ascontiguousarray(symbols, np.uint32)
    freqs = np.ascontiguousarray(freqs, np.uint32)
    first_pos = np.ascontiguousarray(first_pos, np.uint32)
    n = symbols.size
    codes = np.zeros(max(n, 1), np.uint32)
    lens = np.zeros(max(n, 1), np.uint8)
    cap = 64 + 8 * (n + 2) + 2 * 70
    header = np.zeros(cap, np.uint8)
    m, eofc, eofl, hb, pb = C.c_uint32(), C.c_uint32(), C.c_uint8(), C.c_uint64(), C.c_uint64()
    _lib.check(_lib.load().mnenc_debug_plan_sequence(_p(symbols), _p(freqs), _p(first_pos), n, C.byref(m), _p(codes), _p(lens),
                                                     C.byref(eofc), C.byref(eofl), _p(header), cap, C.byref(hb), C.byref(pb)))
    return dict(golomb_m=m.value, codes=codes[:n], lens=lens[:n], eof_code=eofc.value, eof_len=eofl.value,
                header=header[: (hb.value + 7) // 8].tobytes(), header_bits=hb.value, payload_bits=pb.value)
