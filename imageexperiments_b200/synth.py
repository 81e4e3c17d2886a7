"""Synthetic RGB test images with photo-like statistics (SURVEY.md §8d's recipe, numpy RNG).

Low-frequency sinusoids per channel, random half-plane edges and a little Gaussian noise. The same
generated buffer is always fed to the reference, the oracle and the GPU encoder, so only its
statistics (≈4.4 matching-pursuit iterations per block-channel at quality 3.5) matter.
"""
from __future__ import annotations

import numpy as np


def synthetic_rgb(width: int, height: int, seed: int = 12345) -> np.ndarray:
    rng = np.random.default_rng(seed)
    ys, xs = np.meshgrid(np.arange(height, dtype=np.float32), np.arange(width, dtype=np.float32), indexing="ij")
    img = np.full((height, width, 3), 128.0, np.float32)
    for c in range(3):
        for k in range(4):
            fx, fy = rng.random(2) * 0.05 * (k + 1)
            ph = rng.random() * 6.28
            img[..., c] += (40.0 / (k + 1)) * np.sin(fx * xs + fy * ys + ph)
    for _ in range(24):
        x0, y0 = rng.random() * width, rng.random() * height
        ang = rng.random() * 6.28
        side = ((xs - x0) * np.cos(ang) + (ys - y0) * np.sin(ang)) > 0
        for c in range(3):
            img[..., c] += np.where(side, (rng.random() * 2 - 1) * 15.0, 0.0)
    img += rng.normal(0.0, 2.0, img.shape).astype(np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)
