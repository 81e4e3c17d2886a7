"""Small workload for compute-sanitizer (memcheck / racecheck / synccheck): one ragged image through the single-image,
batch and (thread-rank) stripe paths, checked against each other.

    compute-sanitizer --tool memcheck python tools/sanitize_small.py
"""
import os
import sys
import threading

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import imageexperiments_b200 as mn  # noqa: E402
from imageexperiments_b200.synth import synthetic_rgb  # noqa: E402

img = synthetic_rgb(200, 136, 7)
gray = np.full((264, 1608, 3), 77, np.uint8)
ctxs = [mn.createCompressionContextFast(32, 8, 3.5, device=0) for _ in range(3)]
for im in (img, gray):
    want = mn.encodeImageFast(im, ctxs[0])
    assert mn.encodeImageBatch([im, im, im], ctxs[0]) == [want] * 3
    comms = mn.ThreadComm.make(3)
    out = [None] * 3

    def work(r):
        out[r] = mn.encodeImageStripes(im, ctxs[r], comms[r])

    ts = [threading.Thread(target=work, args=(r,)) for r in range(3)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert out[0] == want and out[1] is None
for c in ctxs:
    c.close()
print("sanitize_small OK")
