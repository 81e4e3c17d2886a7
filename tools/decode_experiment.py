import sys, time, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import imageexperiments_b200 as mn
from imageexperiments_b200.synth import synthetic_rgb
W,H=4928,3264
ctx = mn.createCompressionContextFast(32, 8, 3.5, device=0)
imgs=[synthetic_rgb(W,H,12345+i) for i in range(4)]
ring=[mn.encodeImageFast(im, ctx) for im in imgs]
def t(label, hold=False):
    for i in range(3): mn.decodeImageFast(ring[i%4], ctx, zero_copy=True)
    t0=time.perf_counter(); ph=np.zeros(8)
    for i in range(8):
        if hold: img = mn.decodeImageFast(ring[i%4], ctx, zero_copy=True)
        else: mn.decodeImageFast(ring[i%4], ctx, zero_copy=True)
        ph+=ctx.last_timing()[0]
    dt=(time.perf_counter()-t0)/8*1e3
    print(label, round(dt,3), "ms", np.round(ph/8,3).tolist(), flush=True)
t("fresh")
t("fresh, holding previous result", hold=True)
dev=[torch.from_numpy(a).cuda() for a in imgs]
mn.encodeImageBatch([dev[i%4] for i in range(32)], ctx, zero_copy=True)
t("after an encode batch (9 lanes exist)")
out=mn.decodeImageBatch([ring[i%4] for i in range(16)], ctx, zero_copy=True); out=None
t("after a decode batch (12 lanes exist)")
pinned=[]
for a in imgs:
    p=torch.empty(a.shape,dtype=torch.uint8,pin_memory=True); p.numpy()[...]=a; pinned.append(p)
streams=[mn.encodeImageFast(pinned[i], ctx) for i in range(4)]
ring=streams
t("streams from pinned inputs")
