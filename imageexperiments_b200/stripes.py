"""One image, stripe-sharded over the GPUs of a process group (BASELINE.json configs[3]).

    ctx  = createCompressionContextFast(32, 8, 3.5, device=local_rank)
    data = encodeImageStripes(img, ctx, TorchComm())        # rank 0: bytes of the .mn stream, other ranks: None

The reference's encodeImageFast (CompressedImage.cpp:578-633) walks blocks column-major, so every rank takes a contiguous
range of 8-pixel block COLUMNS and owns the matching slice of every symbol stream. Matching pursuit needs no exchange.
The entropy back end runs as the six-phase stripe protocol of include/mnenc.h; the exchanges between the phases are
done here with torch.distributed (NCCL over NVLink on the GPUs, gloo in the CPU tests of the host logic):

    all-gather  stream sizes + last raw DC value      (DC delta chain, CompressedImage.cpp:487-505)
    all-gather  boundary run records                  (run stitching of huffman::runLengthEncode, Huffman.cpp:246-279)
    all-reduce  symbol histogram (SUM), first occurrence (MIN), largest symbol (MAX)   -> identical code tables everywhere
    all-gather  payload bits per (rank, sequence)     (offset scan: where each rank's segments start in the file)
    reduce      packed words to rank 0 (SUM of disjoint bits = concatenation of the per-GPU bitstream segments)
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import List, Optional, Tuple

import numpy as np

from . import _lib
from .compressed import CompressionContextFast, MnBuffer, _take

_vp = C.c_void_p


def stripeColumns(nbx: int, world: int) -> List[Tuple[int, int]]:
    """Block-column range [bx0, bx1) of every rank: contiguous, as even as possible (ranks may be empty when nbx < world)."""
    return [(nbx * r // world, nbx * (r + 1) // world) for r in range(world)]


class _DevArray:
    """Raw device memory of the encoder library as a __cuda_array_interface__ object (int32 view)."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (ptr, False), "version": 2}


def _wrap(ptr: int, n: int, device: int):
    import torch
    return torch.as_tensor(_DevArray(ptr, n), device=torch.device("cuda", device))


class TorchComm:
    """The exchanges of the stripe protocol over a torch.distributed process group (NCCL on GPUs)."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.backend = dist.get_backend(group)

    def all_gather(self, a: np.ndarray, device: int) -> np.ndarray:
        import torch
        t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1))
        if self.backend == "nccl":
            t = t.cuda(device)
        out = torch.empty(self.world * t.numel(), dtype=torch.uint8, device=t.device)  # flat: gloo accepts only the concatenated form
        self.dist.all_gather_into_tensor(out, t, group=self.group)
        return out.cpu().numpy().view(a.dtype).reshape((self.world,) + a.shape)

    def all_reduce_device(self, ptr: int, n: int, op: str, device: int) -> None:
        import torch
        t = _wrap(ptr, n, device)
        ops = {"sum": self.dist.ReduceOp.SUM, "min": self.dist.ReduceOp.MIN, "max": self.dist.ReduceOp.MAX}
        self.dist.all_reduce(t, op=ops[op], group=self.group)
        torch.cuda.current_stream(device).synchronize()

    def reduce_device(self, ptr: int, n: int, device: int) -> None:
        import torch
        t = _wrap(ptr, n, device)
        self.dist.reduce(t, dst=self.dist.get_global_rank(self.group, 0) if self.group is not None else 0, op=self.dist.ReduceOp.SUM,
                         group=self.group)
        torch.cuda.current_stream(device).synchronize()


class ThreadComm:
    """Same exchanges between `world` threads of ONE process (each with its own context, possibly all on one GPU): lets a
    single-GPU box run the whole stripe protocol. Create with ThreadComm.make(world) -> list of per-rank comms."""

    class _Shared:
        def __init__(self, world):
            self.world = world
            self.barrier = threading.Barrier(world)
            self.slots = [None] * world

    def __init__(self, shared, rank):
        self.s, self.rank, self.world = shared, rank, shared.world

    @staticmethod
    def make(world: int) -> List["ThreadComm"]:
        sh = ThreadComm._Shared(world)
        return [ThreadComm(sh, r) for r in range(world)]

    def all_gather(self, a: np.ndarray, device: int) -> np.ndarray:
        self.s.slots[self.rank] = np.ascontiguousarray(a).copy()
        self.s.barrier.wait()
        out = np.stack(self.s.slots)
        self.s.barrier.wait()
        return out

    def _device_op(self, ptr, n, device, fn):
        import torch
        self.s.slots[self.rank] = (ptr, n, device)
        self.s.barrier.wait()
        if self.rank == 0:
            ts = [_wrap(p, m, d) for (p, m, d) in self.s.slots]
            fn(ts)
            torch.cuda.synchronize()
        self.s.barrier.wait()

    def all_reduce_device(self, ptr: int, n: int, op: str, device: int) -> None:
        import torch

        def fn(ts):
            acc = ts[0].clone()
            for t in ts[1:]:
                t = t.to(acc.device)
                acc = acc + t if op == "sum" else (torch.minimum(acc, t) if op == "min" else torch.maximum(acc, t))
            for t in ts:
                t.copy_(acc)
        self._device_op(ptr, n, device, fn)

    def reduce_device(self, ptr: int, n: int, device: int) -> None:
        def fn(ts):
            acc = ts[0].clone()
            for t in ts[1:]:
                acc += t.to(acc.device)
            ts[0].copy_(acc)
        self._device_op(ptr, n, device, fn)


def encodeImageStripes(imgIn, context: CompressionContextFast, comm, width: Optional[int] = None,
                       timings: Optional[dict] = None, zero_copy: bool = False):
    """compressed::encodeImageFast (CompressedImage.cpp:578-633) of ONE image over all ranks of ``comm``.

    ``imgIn``: the full HxWx3 uint8 image in host memory (numpy array or CPU torch tensor; every rank copies only its own
    block columns to its GPU), or — with ``width`` = full image width — a CUDA uint8 tensor [H, stripe width, 3] holding
    just this rank's stripe (columns 8*bx0 .. min(8*bx1, width)). Returns the `.mn` bytes on rank 0, None elsewhere.
    ``zero_copy``: return an MnBuffer over the library's own memory instead of copying it into ``bytes``.
    ``timings`` (optional dict) accumulates host wall seconds per phase / exchange of the protocol."""
    import time
    tl = [time.perf_counter()]

    def lap(name):
        if timings is not None:
            t = time.perf_counter()
            timings[name] = timings.get(name, 0.0) + t - tl[0]
            tl[0] = t

    L = _lib.load()
    K = context.K
    rank, world, dev = comm.rank, comm.world, context.device
    is_np = isinstance(imgIn, np.ndarray)
    on_device = (not is_np) and imgIn.is_cuda
    if is_np:
        if imgIn.dtype != np.uint8 or imgIn.ndim != 3 or imgIn.shape[2] != 3:
            raise ValueError("expected HxWx3 uint8")
        if imgIn.strides[2] != 1 or imgIn.strides[1] != 3:
            imgIn = np.ascontiguousarray(imgIn)
        h, w_in = imgIn.shape[:2]
        base, stride = imgIn.ctypes.data, imgIn.strides[0]
    else:
        if str(imgIn.dtype) != "torch.uint8" or imgIn.dim() != 3 or imgIn.shape[2] != 3 or not imgIn.is_contiguous():
            raise ValueError("expected contiguous HxWx3 uint8")
        h, w_in = int(imgIn.shape[0]), int(imgIn.shape[1])
        base, stride = imgIn.data_ptr(), 3 * w_in
    w = int(width) if on_device else w_in
    if on_device and width is None:
        raise ValueError("a device stripe needs the full image width")
    bx0, bx1 = stripeColumns((w + 7) // 8, world)[rank]
    if on_device:
        if w_in != min(8 * bx1, w) - 8 * bx0:
            raise ValueError("device tensor is not this rank's stripe")
        ptr = base
    else:
        ptr = base + 3 * 8 * bx0

    rec1 = np.zeros(3 * K + 3, np.uint32)
    _lib.check(L.mnenc_stripe_begin(context._h, _vp(ptr), int(on_device), w, h, stride, rank, world, bx0, bx1, rec1.ctypes.data_as(_vp)))
    lap("begin")
    all1 = comm.all_gather(rec1, dev)
    lap("x_gather1")
    rec2 = np.zeros((6 * K, 6), np.uint32)
    _lib.check(L.mnenc_stripe_rle(context._h, all1.ctypes.data_as(_vp), rec2.ctypes.data_as(_vp)))
    lap("rle")
    all2 = comm.all_gather(rec2, dev)
    lap("x_gather2")
    hist, first, msym, nh, nm = _vp(), _vp(), _vp(), C.c_size_t(), C.c_size_t()
    _lib.check(L.mnenc_stripe_histogram(context._h, all2.ctypes.data_as(_vp), C.byref(hist), C.byref(first), C.byref(nh), C.byref(msym),
                                        C.byref(nm)))
    lap("histogram")
    if world > 1:
        comm.all_reduce_device(msym.value, nm.value, "max", dev)
        comm.all_reduce_device(hist.value, nh.value, "sum", dev)
        comm.all_reduce_device(first.value, nh.value, "min", dev)
    lap("x_allreduce")
    bits = np.zeros(6 * K + 1, np.uint64)
    _lib.check(L.mnenc_stripe_plan(context._h, bits.ctypes.data_as(_vp)))
    lap("plan")
    allb = comm.all_gather(bits, dev)
    lap("x_gather3")
    outp, nw, nb = _vp(), C.c_size_t(), C.c_size_t()
    _lib.check(L.mnenc_stripe_pack(context._h, allb.ctypes.data_as(_vp), C.byref(outp), C.byref(nw), C.byref(nb)))
    lap("pack")
    if world > 1:
        comm.reduce_device(outp.value, nw.value, dev)
    lap("x_reduce")
    out, n = _vp(), C.c_size_t()
    _lib.check(L.mnenc_stripe_finish(context._h, C.byref(out), C.byref(n)))
    lap("finish")
    if rank != 0:
        return None
    return MnBuffer(out, n.value) if zero_copy else _take(out, n.value)


# ---- exchanges inside the library (NCCL, csrc/stripe_comm.cu) ----------------------------------------------------------
STRIPE_PHASES = ("begin", "x_sizes", "rle", "x_runs", "histogram", "x_histogram", "tables", "x_offsets", "pack", "x_segments", "finish")


def initStripeComm(context: CompressionContextFast, group=None) -> Tuple[int, int]:
    """Gives ``context`` an NCCL communicator of its own (mnenc_comm_init_rank) over the ranks of a torch.distributed group:
    the 128-byte NCCL id is made on rank 0 and broadcast through the group - the only use of torch here. Without an
    initialised process group the communicator has this one rank (the same code path, N = 1). Returns (rank, world)."""
    import torch
    import torch.distributed as dist
    L = _lib.load()
    ident = np.zeros(128, np.uint8)
    if not (dist.is_available() and dist.is_initialized()):
        rank, world = 0, 1
        _lib.check(L.mnenc_comm_unique_id(ident.ctypes.data_as(_vp)))
    else:
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        if rank == 0:
            _lib.check(L.mnenc_comm_unique_id(ident.ctypes.data_as(_vp)))
        t = torch.from_numpy(ident)
        if dist.get_backend(group) == "nccl":
            t = t.cuda(context.device)
        dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        ident = t.cpu().numpy()
    _lib.check(L.mnenc_comm_init_rank(context._h, rank, world, ident.ctypes.data_as(_vp)))
    context._stripe_rank, context._stripe_world = rank, world
    return rank, world


def encodeImageStriped(imgIn, context: CompressionContextFast, width: Optional[int] = None, timings: Optional[dict] = None,
                       zero_copy: bool = False):
    """compressed::encodeImageFast (CompressedImage.cpp:578-633) of ONE image over all ranks of the context's communicator
    (initStripeComm), every exchange done by the library itself over NCCL (mnenc_encode_striped_rgb8). Arguments as
    encodeImageStripes. Returns the `.mn` bytes on rank 0, None elsewhere."""
    L = _lib.load()
    rank, world = context._stripe_rank, context._stripe_world
    is_np = isinstance(imgIn, np.ndarray)
    on_device = (not is_np) and imgIn.is_cuda
    if is_np:
        if imgIn.dtype != np.uint8 or imgIn.ndim != 3 or imgIn.shape[2] != 3:
            raise ValueError("expected HxWx3 uint8")
        if imgIn.strides[2] != 1 or imgIn.strides[1] != 3:
            imgIn = np.ascontiguousarray(imgIn)
        h, w_in = imgIn.shape[:2]
        base, stride = imgIn.ctypes.data, imgIn.strides[0]
    else:
        if str(imgIn.dtype) != "torch.uint8" or imgIn.dim() != 3 or imgIn.shape[2] != 3 or not imgIn.is_contiguous():
            raise ValueError("expected contiguous HxWx3 uint8")
        h, w_in = int(imgIn.shape[0]), int(imgIn.shape[1])
        base, stride = imgIn.data_ptr(), 3 * w_in
    if on_device and width is None:
        raise ValueError("a device stripe needs the full image width")
    w = int(width) if width is not None else w_in
    bx0, bx1 = stripeColumns((w + 7) // 8, world)[rank]
    stripe_only = width is not None  # the tensor holds just this rank's block columns
    if stripe_only and w_in != min(8 * bx1, w) - 8 * bx0:
        raise ValueError("tensor is not this rank's stripe")
    ptr = base if stripe_only else base + 3 * 8 * bx0
    out, n = _vp(), C.c_size_t()
    _lib.check(L.mnenc_encode_striped_rgb8(context._h, _vp(ptr), int(on_device), w, h, stride, bx0, bx1, C.byref(out), C.byref(n)))
    if timings is not None:
        ms = np.zeros(16, np.float32)
        _lib.check(L.mnenc_last_stripe_timing(context._h, ms.ctypes.data_as(_vp)))
        for name, v in zip(STRIPE_PHASES, ms):
            timings[name] = timings.get(name, 0.0) + float(v) / 1e3
    if rank != 0:
        return None
    return MnBuffer(out, n.value) if zero_copy else _take(out, n.value)
