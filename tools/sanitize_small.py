"""A small tour of the library for compute-sanitizer (memcheck / racecheck): one 200x136 encode (ragged edge: zero padding),
its decode, and the no-allocation scatter entry point. The results are checked against each other, so a run that passes
under the tool is also a correct run. (The striped encode is left out: NCCL under the sanitizer takes minutes.)

    compute-sanitizer --tool memcheck  python tools/sanitize_small.py
    compute-sanitizer --tool racecheck python tools/sanitize_small.py
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import imageexperiments_b200 as mn  # noqa: E402
from imageexperiments_b200 import _lib  # noqa: E402
from imageexperiments_b200.synth import synthetic_rgb  # noqa: E402

K = 32
ctx = mn.createCompressionContextFast(K, 8, 3.5, device=0)
img = synthetic_rgb(200, 136, 7)
data = mn.encodeImageFast(img, ctx)
counts, pairs = mn.lastSymbols(ctx, 200, 136)
pixels = mn.decodeImageFast(data, ctx)
assert pixels.shape == img.shape
parsed = mn.parseStream(data)
L = _lib.load()
d_counts = torch.from_numpy(counts.view(np.int16).reshape(-1)).cuda()
d_pairs = torch.from_numpy(pairs.view(np.int16).reshape(-1)).cuda()
need = 2 * int(counts.astype(np.int64).sum())
d_streams = torch.zeros(need, dtype=torch.int16, device="cuda")
offs = np.zeros(6 * K + 1, np.uint64)
_lib.check(L.mnenc_scatter_streams_static(C.c_void_p(d_counts.data_ptr()), C.c_void_p(d_pairs.data_ptr()), K, counts.shape[0],
                                          C.c_void_p(d_streams.data_ptr()), need, offs.ctypes.data_as(C.c_void_p), None))
got = d_streams.cpu().numpy().view(np.uint16)
o = offs.astype(np.int64)
for s, want in enumerate(parsed["streams"]):
    assert (got[o[s]:o[s + 1]] == want).all(), s
ctx.close()
print(f"sanitize_small OK: {len(data)} bytes, {need} symbols", flush=True)
