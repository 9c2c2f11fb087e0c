#!/usr/bin/env python3
"""A synthetic albedo map for LayeredSkin's "texture albedo" (the fork ships no image files): 64 x 32 RGB texels in (0.2, 1], a
warm base with darker blotches and a few sharp freckles, written as a little-endian PFM (scenes/albedo64.pfm)."""
import os
import sys

import numpy as np


def albedomap(w=64, h=32):
    v, u = np.meshgrid((np.arange(h) + 0.5) / h, (np.arange(w) + 0.5) / w, indexing="ij")
    base = np.stack([0.95 - 0.1 * v, 0.8 - 0.15 * np.sin(3 * np.pi * u) ** 2, 0.7 - 0.2 * v * u], axis=-1)
    blot = 0.25 * (np.sin(9 * u + 4 * v) * np.cos(7 * v - 2 * u)) ** 2
    img = base * (1.0 - blot[..., None])
    rng = np.random.RandomState(3)
    for _ in range(40):
        y, x = rng.randint(0, h), rng.randint(0, w)
        img[y, x] *= (0.45, 0.35, 0.3)
    return np.clip(img, 0.2, 1.0).astype(np.float32)


def write_pfm(path, img):
    h, w, _ = img.shape
    with open(path, "wb") as f:
        f.write(b"PF\n%d %d\n-1.0\n" % (w, h))
        f.write(np.ascontiguousarray(img, "<f4").tobytes())


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes", "albedo64.pfm")
    write_pfm(out, albedomap())
    print(out)
