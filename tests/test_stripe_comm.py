"""CPU-only, world_size 2 over gloo: the host side of the stripe protocol's exchanges (imageexperiments_b200.stripes.TorchComm)
and the column split. The device-side collectives need GPUs (tests/test_gpu_stripes.py, tools/stripe_check.py)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from imageexperiments_b200.stripes import TorchComm, stripeColumns
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    comm = TorchComm()
    assert (comm.rank, comm.world, comm.backend) == (rank, world, "gloo")
    # the three record shapes of the protocol: u32 [3K+3], u32 [6K][6], u64 [6K+1]
    K = 32
    rec1 = (np.arange(3 * K + 3, dtype=np.uint32) + 1000 * rank)
    rec2 = (np.arange(6 * K * 6, dtype=np.uint32).reshape(6 * K, 6) * (rank + 1))
    bits = (np.arange(6 * K + 1, dtype=np.uint64) << np.uint64(33)) + np.uint64(rank)
    ok = True
    for a in (rec1, rec2, bits):
        g = comm.all_gather(a, device=0)
        ok &= g.shape == (world,) + a.shape and g.dtype == a.dtype and np.array_equal(g[rank], a)
        other = 1 - rank
        if a is rec1:
            ok &= np.array_equal(g[other], np.arange(3 * K + 3, dtype=np.uint32) + 1000 * other)
        elif a is rec2:
            ok &= np.array_equal(g[other], np.arange(6 * K * 6, dtype=np.uint32).reshape(6 * K, 6) * (other + 1))
        else:
            ok &= np.array_equal(g[other], (np.arange(6 * K + 1, dtype=np.uint64) << np.uint64(33)) + np.uint64(other))
    cols = stripeColumns(616, world)
    ok &= cols[0][0] == 0 and cols[-1][1] == 616 and all(cols[i][1] == cols[i + 1][0] for i in range(world - 1))
    q.put((rank, bool(ok)))
    dist.barrier()
    dist.destroy_process_group()


def test_torchcomm_all_gather_over_gloo():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert got == [(0, True), (1, True)]


@pytest.mark.parametrize("nbx,world", [(616, 8), (2048, 8), (3, 8), (1, 2), (17, 5)])
def test_stripe_columns_partition(nbx, world):
    from imageexperiments_b200.stripes import stripeColumns
    cols = stripeColumns(nbx, world)
    assert len(cols) == world and cols[0][0] == 0 and cols[-1][1] == nbx
    assert all(a[1] == b[0] for a, b in zip(cols, cols[1:]))
    sizes = [b - a for a, b in cols]
    assert max(sizes) - min(sizes) <= 1
