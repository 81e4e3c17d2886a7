"""Batch mode on the GPU box: frames that share one matching pursuit (mnenc_encode_batch_rgb8, frame groups) must give the bytes
of the one-at-a-time encode; then the 1080p throughput (BASELINE.json configs[4]).   python tools/batch_check.py"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import imageexperiments_b200 as mn
from imageexperiments_b200.synth import synthetic_rgb

ctx = mn.createCompressionContextFast(32, 8, 3.5, device=0)
bad = 0
for (w, h, n) in ([] if os.environ.get("BATCH_SKIP_PARITY") else [(100, 52, 5), (8, 8, 3), (1023, 77, 7), (640, 360, 33), (1920, 1080, 17)]):
    frames = [synthetic_rgb(w, h, 100 + i) for i in range(n)]
    want = [bytes(mn.encodeImageFast(f, ctx)) for f in frames]
    got_host = [bytes(b) for b in mn.encodeImageBatch(frames, ctx)]
    got_dev = [bytes(b) for b in mn.encodeImageBatch([torch.from_numpy(f).cuda() for f in frames], ctx)]
    ok = got_host == want and got_dev == want
    bad += 0 if ok else 1
    print(f"{n} frames of {w}x{h}: host batch identical {got_host == want}, device batch identical {got_dev == want}", flush=True)
fw, fh, nf = 1920, 1080, int(os.environ.get("BATCH_FRAMES", "64"))
distinct = [torch.from_numpy(synthetic_rgb(fw, fh, 12345 + f)).cuda() for f in range(8)]
frames = [distinct[i % 8] for i in range(nf)]
for name in ("resident",):
    mn.encodeImageBatch(frames, ctx, zero_copy=True)
    t0 = time.perf_counter()
    for _ in range(3):
        mn.encodeImageBatch(frames, ctx, zero_copy=True)
    dt = (time.perf_counter() - t0) / 3
    ms, launches = ctx.last_timing()
    print(f"1080p x{nf} {name}: {nf / dt:.1f} frames/s, {nf * fw * fh / 1e6 / dt:.1f} MP/s; mp stage sum {ms[1]:.2f} ms, launches {launches}", flush=True)
ctx.close()
sys.exit(1 if bad else 0)
