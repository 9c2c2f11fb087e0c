#!/usr/bin/env python3
"""Synthetic stand-ins, in the file format the head scenes' own textures use (lambertian-norm2.tga, bump-lowRes2.tga: not shipped
with the fork), with sizes that are NOT powers of two so that MIPMap's Lanczos zoom runs:
  scenes/albedo48x40.tga   true colour, 24 bit, run-length encoded (type 10), origin top-left
  scenes/bump48x40.tga     grey, 8 bit, raw (type 3), origin bottom-left
tests/test_tga_textures.py compares the host mirror's pyramids of both with the real reference's (tests/golden/tex_tga.npz)."""
import os
import struct

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H = 48, 40


def header(image_type, depth, descriptor):
    return struct.pack("<BBBHHBHHHHBB", 0, 0, image_type, 0, 0, 0, 0, 0, W, H, depth, descriptor)


def rle_rows(rows):
    out = bytearray()
    for row in rows:                       # packets do not cross scan lines
        x = 0
        while x < len(row):
            run = 1
            while x + run < len(row) and run < 128 and bytes(row[x + run]) == bytes(row[x]):
                run += 1
            if run > 1:
                out.append(0x80 | (run - 1)); out += bytes(row[x]); x += run
            else:
                lit = 1
                while x + lit < len(row) and lit < 128 and not (x + lit + 1 < len(row) and bytes(row[x + lit]) == bytes(row[x + lit + 1])):
                    lit += 1
                out.append(lit - 1)
                for k in range(lit):
                    out += bytes(row[x + k])
                x += lit
    return bytes(out)


def main():
    yy, xx = np.mgrid[0:H, 0:W]
    u, v = (xx + 0.5) / W, (yy + 0.5) / H
    r = 0.55 + 0.35 * np.sin(2 * np.pi * (u + 0.3 * v))
    g = 0.45 + 0.30 * np.cos(2 * np.pi * (2 * v - u))
    b = 0.35 + 0.25 * np.sin(2 * np.pi * 3 * u) * np.cos(2 * np.pi * v)
    rgb = np.clip(np.stack([b, g, r], axis=-1) * 255, 0, 255).astype(np.uint8)      # BGR order in the file
    rgb[5:9, 10:30] = (40, 90, 200)                                                # a flat patch: real runs for the RLE packets
    with open(os.path.join(ROOT, "scenes", "albedo48x40.tga"), "wb") as f:
        f.write(header(10, 24, 0x20))
        f.write(rle_rows(rgb))
    grey = np.clip(127.5 + 100 * np.sin(2 * np.pi * 4 * u) * np.sin(2 * np.pi * 3 * v), 0, 255).astype(np.uint8)
    with open(os.path.join(ROOT, "scenes", "bump48x40.tga"), "wb") as f:
        f.write(header(3, 8, 0x00))
        f.write(grey.tobytes())
    # second pair, the remaining encodings: true colour with an alpha byte (32 bit), raw, bottom-left origin, stored right to left;
    # grey, run-length encoded (type 11), top-left origin
    bgra = np.concatenate([rgb, np.full((H, W, 1), 255, np.uint8)], axis=-1)
    with open(os.path.join(ROOT, "scenes", "albedo48x40_b.tga"), "wb") as f:
        f.write(header(2, 32, 0x18))                   # 8 alpha bits + right-to-left
        f.write(bgra[:, ::-1].tobytes())
    grey2 = grey.copy(); grey2[12:20, 4:40] = 200
    with open(os.path.join(ROOT, "scenes", "bump48x40_b.tga"), "wb") as f:
        f.write(header(11, 8, 0x20))
        f.write(rle_rows(grey2[:, :, None]))
    print("wrote scenes/albedo48x40.tga, scenes/bump48x40.tga, scenes/albedo48x40_b.tga, scenes/bump48x40_b.tga")


if __name__ == "__main__":
    main()
