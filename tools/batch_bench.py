"""Throughput of the batch mode (mnenc_encode_batch_rgb8) against one-at-a-time encodes, per lane count.

    python tools/batch_bench.py [W H n_images]        # MNENC_LANES is swept inside (fresh context per setting)
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

    import imageexperiments_b200 as mn
    from imageexperiments_b200.synth import synthetic_rgb

    w = int(sys.argv[1]) if len(sys.argv) > 1 else 4928
    h = int(sys.argv[2]) if len(sys.argv) > 2 else 3264
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 16
    base = synthetic_rgb(w, h, 12345)
    distinct = min(n, 4)
    imgs = [np.ascontiguousarray(np.roll(base, (8 * 37 * i, 8 * 53 * i), axis=(0, 1))) for i in range(distinct)]
    dev = [torch.from_numpy(a).cuda() for a in imgs]
    pin = []
    for a in imgs:
        p = torch.empty(a.shape, dtype=torch.uint8, pin_memory=True)
        p.numpy()[...] = a
        pin.append(p)
    ref = None
    res = {"workload": f"{n} x {w}x{h} RGB synthetic, quality 3.5"}
    for lanes in [int(x) for x in os.environ.get("LANES_SWEEP", "1,2,3,4").split(",")]:
        os.environ["MNENC_LANES"] = str(lanes)
        ctx = mn.createCompressionContextFast(32, 8, 3.5, device=0)
        for name, src in (("resident", dev), ("pinned", pin)):
            batch = [src[i % distinct] for i in range(n)]
            mn.encodeImageBatch(batch, ctx)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(3):
                out = mn.encodeImageBatch(batch, ctx, zero_copy=True)
            dt = (time.perf_counter() - t0) / 3
            res[f"lanes{lanes}_{name}_mp_s"] = round(n * w * h / 1e6 / dt, 1)
        if ref is None:
            ref = [mn.encodeImageFast(imgs[i], ctx) for i in range(distinct)]
        assert all(out[i] == ref[i % distinct] for i in range(n)), "batch stream differs from the single-image stream"
        ctx.close()
    print(json.dumps(res))


if __name__ == "__main__":
    main()
