"""Matching-pursuit check on the GPU box: symbols of the CUDA path against the CPU oracle on seeded images of several sizes and
qualities (first mismatches printed), then the kernel time on the 4928x3264 bench image. Development aid; the parity tests
proper live in tests/test_gpu_parity.py.

    python tools/mp_check.py [--sizes 64x64,200x136,512x512] [--q 3.5,8] [--time 1]
"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import imageexperiments_b200 as mn
from imageexperiments_b200.synth import synthetic_rgb
from oracle.bindings import Dictionary as ODict, OracleLib

ap = argparse.ArgumentParser()
ap.add_argument("--sizes", default="64x64,200x136,512x512")
ap.add_argument("--q", default="3.5,8")
ap.add_argument("--time", type=int, default=1)
a = ap.parse_args()

bad = 0
for q in [float(x) for x in a.q.split(",")]:
    ctx = mn.createCompressionContextFast(32, 8, q, device=0)
    d = ctx.dictionary
    oc = OracleLib().create(ODict(32, 8, d.base, d.detail_rows, d.detail, ctx.Quant))
    for sz in a.sizes.split(","):
        w, h = (int(x) for x in sz.split("x"))
        img = synthetic_rgb(w, h, seed=7)
        data = mn.encodeImageFast(img, ctx)
        counts, pairs = mn.lastSymbols(ctx, w, h)
        oc_counts, oc_pairs = oc.symbols(img)
        same_c = counts == oc_counts
        same_p = (pairs == oc_pairs).all(axis=(2, 3)) if pairs.ndim == 4 else (pairs == oc_pairs)
        nbad = int((~(same_c & same_p)).sum())
        ms, launches = ctx.last_timing()
        print(f"q={q} {w}x{h}: block-channels {counts.size}, differing {nbad}, bytes {len(data)}, oracle-bytes-equal "
              f"{data == oc.encode(img)}, launches {launches}, mp {ms[1]:.3f} ms", flush=True)
        if nbad:
            bad += nbad
            idx = np.argwhere(~(same_c & same_p))[:6]
            for b, ch in idx:
                print(f"   blk {b} ch {ch}: count {counts[b, ch]} vs {oc_counts[b, ch]}; gpu {pairs[b, ch, :6].tolist()} oracle "
                      f"{oc_pairs[b, ch, :6].tolist()}", flush=True)
    ctx.close()

if a.time:
    ctx = mn.createCompressionContextFast(32, 8, 3.5, device=0)
    img = torch.from_numpy(synthetic_rgb(4928, 3264, 12345)).cuda()
    best = None
    for i in range(6):
        t0 = time.perf_counter()
        data = mn.encodeImageFast(img, ctx)
        wall = (time.perf_counter() - t0) * 1e3
        ms, launches = ctx.last_timing()
        best = ms[1] if best is None else min(best, ms[1])
        print(f"16MP encode {i}: wall {wall:.2f} ms, stages {[round(float(x), 3) for x in ms]}, launches {launches}, bytes {len(data)}", flush=True)
    flops, nsym, nbits = ctx.last_work()
    print(f"16MP: mp best {best:.3f} ms, {flops / 1e9:.2f} GFLOP -> {flops / best / 1e9:.1f} TFLOP/s", flush=True)
sys.exit(1 if bad else 0)
