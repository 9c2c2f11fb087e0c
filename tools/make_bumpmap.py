#!/usr/bin/env python3
"""A synthetic displacement map for LayeredSkin's "texture bumpmap" (the fork ships no image files): 64 x 64 one-channel texels,
smooth pores and wrinkles of +-0.006 scene units, written as a little-endian PFM (scenes/bump64.pfm)."""
import os
import sys

import numpy as np


def bumpmap(n=64):
    v, u = np.meshgrid((np.arange(n) + 0.5) / n, (np.arange(n) + 0.5) / n, indexing="ij")
    h = 0.004 * np.sin(2 * np.pi * 7 * u) * np.sin(2 * np.pi * 5 * v) + 0.002 * np.sin(2 * np.pi * (11 * u + 3 * v))
    return h.astype(np.float32)


def write_pfm1(path, img):
    h, w = img.shape
    with open(path, "wb") as f:
        f.write(b"Pf\n%d %d\n-1.0\n" % (w, h))
        f.write(np.ascontiguousarray(img, "<f4").tobytes())


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes", "bump64.pfm")
    write_pfm1(out, bumpmap())
    print(out)
