#!/usr/bin/env python3
"""A synthetic alpha mask for "texture alpha" on a trianglemesh (the fork ships no image files): 32 x 32 luminance texels, 1
with round holes of exact zeros and soft rims, written as a one-channel little-endian PFM (scenes/alpha32.pfm)."""
import os
import sys

import numpy as np


def alphamap(n=32):
    v, u = np.meshgrid((np.arange(n) + 0.5) / n, (np.arange(n) + 0.5) / n, indexing="ij")
    a = np.ones((n, n), np.float32)
    for cu, cv, r in ((0.3, 0.35, 0.13), (0.7, 0.6, 0.16), (0.5, 0.85, 0.1), (0.15, 0.75, 0.08)):
        d = np.hypot(u - cu, v - cv)
        a = np.minimum(a, np.clip((d - r) / 0.05, 0.0, 1.0).astype(np.float32))      # 0 inside, a soft rim of two texels
    return a


def write_pfm1(path, img):
    h, w = img.shape
    with open(path, "wb") as f:
        f.write(b"Pf\n%d %d\n-1.0\n" % (w, h))
        f.write(np.ascontiguousarray(img, "<f4").tobytes())


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes", "alpha32.pfm")
    write_pfm1(out, alphamap())
    print(out)
