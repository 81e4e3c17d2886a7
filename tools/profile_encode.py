"""Short driver for ncu captures: a few full-size encodes with device-resident pixels."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import imageexperiments_b200 as mn
from imageexperiments_b200.synth import synthetic_rgb
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
q = float(sys.argv[2]) if len(sys.argv) > 2 else 3.5
ctx = mn.createCompressionContextFast(32, 8, q, device=0)
img = torch.from_numpy(synthetic_rgb(4928, 3264, 12345)).cuda()
for i in range(n):
    data = mn.encodeImageFast(img, ctx)
ms, launches = ctx.last_timing()
print(len(data), [round(float(x), 3) for x in ms], launches)
