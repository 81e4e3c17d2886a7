"""One process, several GPUs (mnenc_create_multi): encodeImageFast on the multi-GPU context shards the image by block columns over
the GPUs and must return the bytes of a single-GPU context.

    python tools/multi_check.py [n_gpus]
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import imageexperiments_b200 as mn
from imageexperiments_b200.synth import synthetic_rgb

n = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
one = mn.createCompressionContextFast(32, 8, 3.5, device=0)
multi = mn.createCompressionContextFast(32, 8, 3.5, device=list(range(n)))
for (w, h) in [(8, 8), (200, 136), (1023, 77), (4928, 3264)]:
    img = synthetic_rgb(w, h, 12345)
    want = mn.encodeImageFast(img, one)
    got = mn.encodeImageFast(img, multi)
    t0 = time.perf_counter()
    got = mn.encodeImageFast(img, multi)
    dt = time.perf_counter() - t0
    assert got == want, f"{w}x{h}: multi-GPU context differs from the single-GPU context"
    print(f"{w}x{h}: {len(got)} bytes identical over {n} GPUs of one process, {dt * 1e3:.2f} ms", flush=True)
multi.setQuant(mn.createQuantizationTables(32, 8, 6.0))
one.setQuant(mn.createQuantizationTables(32, 8, 6.0))
img = synthetic_rgb(640, 480, 3)
assert mn.encodeImageFast(img, multi) == mn.encodeImageFast(img, one)
# a failure on one GPU of the group (test hook: GPU 0's histogram phase) comes back as ONE error naming that GPU - no thread is left
# waiting in NCCL - and the group encodes correctly right afterwards
from imageexperiments_b200 import _lib
_lib.check(_lib.load().mnenc_debug_fail_next_backend(multi._h))
try:
    mn.encodeImageFast(img, multi)
except mn.MnencError as e:
    assert e.code == -4 and "GPU 0" in str(e) and "injected failure" in str(e), str(e)
else:
    raise AssertionError("the injected failure was not reported")
assert mn.encodeImageFast(img, multi) == mn.encodeImageFast(img, one)
print("collective failure handling OK (one process)", flush=True)
multi.close()
one.close()
print("multi_check OK", flush=True)
