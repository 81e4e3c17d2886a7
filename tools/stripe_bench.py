"""torchrun script: throughput and per-phase times of the stripe protocol (BASELINE.json configs[3]: one 16384x16384 RGB
image sharded by block columns over the ranks). Stripes are resident in HBM (value) or in pinned host memory (e2e).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/stripe_bench.py [W H steps]
"""
import json
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

    w = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    h = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    world_env = int(os.environ.get("WORLD_SIZE", "1"))
    if world_env > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        comm = mn.TorchComm()
    else:
        comm = mn.ThreadComm.make(1)[0]
    rank, world = comm.rank, comm.world
    ctx = mn.createCompressionContextFast(32, 8, 3.5, device=local)
    bx0, bx1 = mn.stripeColumns((w + 7) // 8, world)[rank]
    x0, x1 = 8 * bx0, min(8 * bx1, w)
    # the rank's stripe of the tiled synthetic image (the full 805 MB image is never materialised on one rank)
    tile = synthetic_rgb(2048, 2048, 12345)
    reps_y = (h + 2047) // 2048
    cols = np.concatenate([tile[:, (x % 2048):(x % 2048) + 1] for x in range(x0, x1)], axis=1) if (x1 - x0) < 2048 else None
    if cols is None:
        xs = np.arange(x0, x1) % 2048
        cols = tile[:, xs]
    stripe = np.ascontiguousarray(np.tile(cols, (reps_y, 1, 1))[:h])
    dev = torch.from_numpy(stripe).cuda()
    pin = torch.empty(stripe.shape, dtype=torch.uint8, pin_memory=True)
    pin.numpy()[...] = stripe

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run(inp, n, timings=None):
        nbytes = 0
        barrier()
        t0 = time.perf_counter()
        for _ in range(n):
            # host stripes are addressed as a full-width image whose first column is the stripe's: pass stripe-only tensors
            got = mn.encodeImageStripes(inp, ctx, comm, width=w, timings=timings, zero_copy=True) if inp.is_cuda else encode_pinned(inp, timings)
            nbytes = len(got) if got is not None else 0
        barrier()
        return (time.perf_counter() - t0) / n, nbytes

    def encode_pinned(p, timings):
        # pinned stripe -> device stripe copy is part of the timed call (e2e)
        d = p.cuda(non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return mn.encodeImageStripes(d, ctx, comm, width=w, timings=timings, zero_copy=True)

    run(dev, 2)
    tm = {}
    dt, nbytes = run(dev, steps, tm)
    dt_e2e, _ = run(pin, steps)
    ms, launches = ctx.last_timing()
    res = torch.tensor([dt, dt_e2e, float(ms[1])], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(res, op=dist.ReduceOp.MAX)
    if rank == 0:
        mpix = w * h / 1e6
        print(json.dumps({
            "workload": f"{w}x{h} RGB synthetic (2048^2 tile repeated), quality 3.5, ONE image stripe-sharded by block columns over {world} GPU(s)",
            "n_gpus": world, "steps": steps, "value_mp_s": round(mpix / float(res[0]), 1), "ms_per_image": round(float(res[0]) * 1e3, 2),
            "e2e_mp_s": round(mpix / float(res[1]), 1), "e2e_ms_per_image": round(float(res[1]) * 1e3, 2),
            "mp_kernel_ms_max_over_ranks": round(float(res[2]), 2), "bytes_out": nbytes,
            "rank0_phase_ms": {k: round(v / steps * 1e3, 3) for k, v in tm.items()}, "launches_rank0": launches}), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
