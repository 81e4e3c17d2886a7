"""Decoder throughput on one B200: BASELINE.json configs[1] geometry (4928 x 3264, q 3.5), the stream
produced by the GPU encoder. Prints one JSON line: end-to-end MP/s (host stream in, host pixels out), with the pixels left
on the device, the phases of one call, and the CPU baselines (reference decoder from oracle/_ref when present, oracle port)
on a bounded sample.   python tools/decode_bench.py [--steps 10] [--cpu 1]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import imageexperiments_b200 as mn  # noqa: E402
from imageexperiments_b200.synth import synthetic_rgb  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cpu", type=int, default=1)
    ap.add_argument("--batch", type=int, default=32, help="streams per decodeImageBatch call (0: skip the batch pass)")
    a = ap.parse_args()
    import torch
    W, H = 4928, 3264
    ctx = mn.createCompressionContextFast(32, 8, 3.5, device=0)
    ring = [mn.encodeImageFast(synthetic_rgb(W, H, 12345 + i), ctx) for i in range(4)]  # different streams: no cache-warm repeats
    mp = W * H / 1e6
    out = torch.empty((H, W, 3), dtype=torch.uint8, device="cuda:0")

    def run(fn):
        for i in range(a.warmup):
            fn(ring[i % 4])
        torch.cuda.synchronize()
        t = time.perf_counter()
        phases = np.zeros(8)
        for i in range(a.steps):
            fn(ring[i % 4])
            phases += ctx.last_timing()[0]
        torch.cuda.synchronize()
        return (time.perf_counter() - t) / a.steps * 1e3, phases / a.steps

    ms_host, ph_host = run(lambda d: mn.decodeImageFast(d, ctx, zero_copy=True))
    ms_dev, ph_dev = run(lambda d: mn.decodeImageFast(d, ctx, out=out))
    # throughput mode: 32 streams per call over the lanes (host threads run the entropy decode of different streams in parallel)
    batch = [ring[i % 4] for i in range(a.batch)]
    imgs = None
    for _ in range(2 if a.batch else 0):
        imgs = None  # hand the 32 pinned blocks back before the next batch asks for 32
        imgs = mn.decodeImageBatch(batch, ctx, zero_copy=True)
    reps = max(1, a.steps // 3) if a.batch else 0
    t = time.perf_counter()
    for _ in range(reps):
        imgs = None
        imgs = mn.decodeImageBatch(batch, ctx, zero_copy=True)
    ms_batch = (time.perf_counter() - t) / max(reps, 1) * 1e3
    imgs = None
    line = {"metric": "decode_throughput", "unit": "MP/s", "workload": f"{W}x{H} q3.5 K32, stream from the GPU encoder ({len(ring[0])} B)",
            "e2e": {"value": round(mp / ms_host * 1e3, 1), "ms": round(ms_host, 3), "h2d_bytes_per_step": len(ring[0]), "d2h_bytes_per_step": 3 * W * H},
            "pixels_on_device": {"value": round(mp / ms_dev * 1e3, 1), "ms": round(ms_dev, 3)},
            "batch_e2e": {"streams": a.batch, "value": round(a.batch * mp / ms_batch * 1e3, 1) if a.batch else None, "ms_per_batch": round(ms_batch, 2),
                          "lanes": int(os.environ.get("MNENC_LANES", 6))},
            "phases_ms": {"lengths_host": round(float(ph_dev[0]), 3), "streams_host": round(float(ph_dev[1]), 3),
                          "upload_kernels_host_clock": round(float(ph_dev[2]), 3), "kernels_gather_reconstruct": round(float(ph_dev[3]), 3)},
            "kernel_roofline": {"bound": "hbm", "algorithmic_bytes": None}, "steps": a.steps, "warmup": a.warmup}
    # algorithmic bytes of the two kernels: symbols read once from the arena and once as pairs (+ written once), counts twice, pixels written
    _, nsym, _ = ctx.last_work()
    nbc = 3 * ((W + 7) // 8) * ((H + 7) // 8)
    n = nsym - nbc
    alg = 2 * n + 2 * (2 * n) + 2 * 2 * nbc + 3 * W * H
    line["kernel_roofline"] = {"bound": "hbm", "algorithmic_bytes": int(alg), "achieved_GBps": round(alg / (float(ph_dev[3]) * 1e-3) / 1e9, 1)}
    if a.cpu:
        from oracle.bindings import OracleLib, RefLib, ref_available
        from oracle.bindings import Dictionary as ODict
        d = ctx.dictionary
        small = synthetic_rgb(512, 512, 12345)
        data = mn.encodeImageFast(small, ctx)
        oc = OracleLib().create(ODict(32, 8, d.base, d.detail_rows, d.detail, np.ones(96, np.float32)))
        t = time.perf_counter()
        want = oc.decode(data)
        t_port = time.perf_counter() - t
        assert (want == mn.decodeImageFast(data, ctx)).all()
        line["cpu_baseline"] = {"kind": "port", "value": round(0.262144 / t_port, 3), "unit": "MP/s", "cores": 1, "sample": "512x512 q3.5"}
        if ref_available():
            R = RefLib()
            t = time.perf_counter()
            ref = R.decode(data)
            t_ref = time.perf_counter() - t
            assert (ref == want).all()
            line["cpu_reference"] = {"kind": "reference", "value": round(0.262144 / t_ref, 3), "unit": "MP/s", "cores": 1,
                                     "sample": "512x512 q3.5 (decodeImageFast incl. its dictionary rebuild)"}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
