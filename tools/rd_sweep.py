"""Rate-distortion sweep of the encode -> decode cycle on one B200: the per-image part of the reference CLI's `-g` mode
(Compression/Compression.cpp:304-378: quality steps, bpp and PSNR of `-n` per step), as a regression report. Inputs:
binary PPM files, or the synthetic generator when none is given. Prints one JSON line per (image, quality).

    python tools/rd_sweep.py [--q 1 2 3 3.5 4 5 6 7 8] [--size 1920x1080] [image.ppm ...]
"""
import argparse
import json
import os
import re
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import imageexperiments_b200 as mn  # noqa: E402
from imageexperiments_b200.synth import synthetic_rgb  # noqa: E402


def read_ppm(path: str) -> np.ndarray:
    raw = open(path, "rb").read()
    m = re.match(rb"P6\s+(?:#[^\n]*\n\s*)*(\d+)\s+(?:#[^\n]*\n\s*)*(\d+)\s+(?:#[^\n]*\n\s*)*255\s", raw)
    if not m:
        raise ValueError(f"{path}: not an 8-bit binary PPM")
    w, h = int(m.group(1)), int(m.group(2))
    return np.frombuffer(raw, np.uint8, 3 * w * h, m.end()).reshape(h, w, 3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--q", type=float, nargs="+", default=[1, 2, 3, 3.5, 4, 5, 6, 7, 8])
    ap.add_argument("--size", default="1920x1080")
    ap.add_argument("images", nargs="*")
    a = ap.parse_args()
    if a.images:
        inputs = [(p, read_ppm(p)) for p in a.images]
    else:
        w, h = (int(x) for x in a.size.split("x"))
        inputs = [(f"synthetic_{w}x{h}_seed12345", synthetic_rgb(w, h, 12345))]
    ctx = mn.createCompressionContextFast(32, 8, 3.5, device=0)
    for name, img in inputs:
        for q in a.q:
            ctx.setQuant(mn.createQuantizationTables(32, 8, q))
            t = time.perf_counter()
            data = mn.encodeImageFast(img, ctx)
            dec = mn.decodeImageFast(data, ctx)
            ms = (time.perf_counter() - t) * 1e3
            print(json.dumps({"image": name, "quality": q, "bpp": round(8 * len(data) / (img.shape[0] * img.shape[1]), 4),
                              "psnr": round(mn.calculatePSNR(img, dec), 3), "bytes": len(data), "cycle_ms": round(ms, 2)}))


if __name__ == "__main__":
    main()
