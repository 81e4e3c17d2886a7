"""torchrun script: the stripe protocol over NCCL, one process per GPU, checked against the single-GPU encode of rank 0 - both with
the exchanges inside the library (mnenc_encode_striped_rgb8, csrc/stripe_comm.cu) and with the caller's transport (TorchComm).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/stripe_check.py [W H]
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    import imageexperiments_b200 as mn
    from imageexperiments_b200.synth import synthetic_rgb

    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    ctx = mn.createCompressionContextFast(32, 8, 3.5, device=local)
    comm = mn.TorchComm()
    mn.initStripeComm(ctx)
    sizes = [(512, 512), (1023, 77), (8, 8), (4928, 3264)]
    if len(sys.argv) >= 3:
        sizes = [(int(sys.argv[1]), int(sys.argv[2]))]
    for (w, h) in sizes:
        img = synthetic_rgb(w, h, 12345) if w * h <= 4928 * 3264 else np.tile(synthetic_rgb(2048, 2048, 12345), ((h + 2047) // 2048, (w + 2047) // 2048, 1))[:h, :w]
        got = mn.encodeImageStripes(img, ctx, comm)
        lib = mn.encodeImageStriped(img, ctx)
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        got = mn.encodeImageStripes(img, ctx, comm)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        dist.barrier()
        tm = {}
        t0 = time.perf_counter()
        lib = mn.encodeImageStriped(img, ctx, timings=tm)
        dl = time.perf_counter() - t0
        if rank == 0:
            t1 = time.perf_counter()
            want = mn.encodeImageFast(img, ctx)
            d1 = time.perf_counter() - t1
            assert got == want, f"{w}x{h}: stripe-sharded stream (caller's transport) differs from the single-GPU stream"
            assert lib == want, f"{w}x{h}: stripe-sharded stream (exchanges inside the library) differs from the single-GPU stream"
            print(f"{w}x{h}: {len(got)} bytes identical over {world} ranks; in-library {dl * 1e3:.2f} ms, caller's transport {dt * 1e3:.2f} ms, "
                  f"single GPU {d1 * 1e3:.2f} ms; phases {({k: round(v * 1e3, 3) for k, v in tm.items()})}", flush=True)
        else:
            assert got is None and lib is None
        dist.barrier()
    if world > 1 and len(sys.argv) < 3:
        check_collective_failures(mn, ctx, rank, world)
    if rank == 0:
        print("stripe_check OK", flush=True)
    ctx.close()
    dist.destroy_process_group()


def check_collective_failures(mn, ctx, rank, world):
    """One rank fails in the middle of a striped encode: nobody may be left waiting inside NCCL. Every rank has to come back with
    an error from the same exchange (the failing rank with its own code, its peers with MNENC_ERR_STATE naming it), and the very
    next striped encode on the same communicator must give the right bytes again."""
    import ctypes as C

    import torch.distributed as dist
    from imageexperiments_b200 import _lib
    from imageexperiments_b200.synth import synthetic_rgb

    L = _lib.load()
    img = synthetic_rgb(640, 512, 99)
    want = mn.encodeImageFast(img, ctx) if rank == 0 else None
    bad = world - 1

    def expect_group_error(own_code, what):
        try:
            what()
        except mn.MnencError as e:
            assert e.code == (own_code if rank == bad else -3), f"rank {rank}: {e}"
            if rank != bad:
                assert f"rank {bad} of the group failed" in str(e), str(e)
        else:
            raise AssertionError(f"rank {rank}: the striped encode succeeded although rank {bad} failed")
        dist.barrier()
        got = mn.encodeImageStriped(img, ctx)  # the group still works
        assert (got == want) if rank == 0 else got is None

    # (1) the histogram phase fails on the last rank (test hook): the code travels with the entry-count gather of the histogram exchange
    if rank == bad:
        _lib.check(L.mnenc_debug_fail_next_backend(ctx._h))
    expect_group_error(-4, lambda: mn.encodeImageStriped(img, ctx))

    # (2) the last rank passes block columns outside the image: mnenc_stripe_begin fails there, the code travels with the first gather
    def bad_columns():
        nbx = 640 // 8
        bx0, bx1 = mn.stripeColumns(nbx, world)[rank]
        if rank == bad:
            bx1 = nbx + 5
        out, n = C.c_void_p(), C.c_size_t()
        _lib.check(L.mnenc_encode_striped_rgb8(ctx._h, C.c_void_p(img.ctypes.data + 24 * bx0), 0, 640, 512, img.strides[0], bx0, bx1, C.byref(out),
                                               C.byref(n)))
    expect_group_error(-1, bad_columns)
    if rank == 0:
        print(f"collective failure handling OK over {world} ranks (histogram phase, begin phase)", flush=True)


if __name__ == "__main__":
    main()
