#!/usr/bin/env python3
"""A small synthetic environment map for LightSource "infinite" (the fork ships none: SURVEY.md F2): 16 x 8 texels (powers of
two, so that MIPMap does not resample), a blue-to-warm vertical gradient with a bright patch, written as a little-endian PFM
(rows in file order = the order ReadImagePFM hands them to the light, core/imageio.cpp:315-395)."""
import os
import sys

import numpy as np


def envmap(w=16, h=8):
    v, u = np.meshgrid((np.arange(h) + 0.5) / h, (np.arange(w) + 0.5) / w, indexing="ij")
    img = np.zeros((h, w, 3), np.float32)
    img[..., 0] = 0.4 + 1.2 * v
    img[..., 1] = 0.6 + 0.5 * v + 0.2 * np.sin(2 * np.pi * u)
    img[..., 2] = 1.4 - 0.9 * v
    img[2, 11] = (28.0, 24.0, 16.0)       # a "sun": most of the importance-sampling mass
    img[3, 11] = (9.0, 8.0, 6.0)
    return img.astype(np.float32)


def write_pfm(path, img):
    h, w, _ = img.shape
    with open(path, "wb") as f:
        f.write(b"PF\n%d %d\n-1.0\n" % (w, h))
        f.write(np.ascontiguousarray(img, "<f4").tobytes())


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes", "env16x8.pfm")
    write_pfm(out, envmap())
    print(out)
