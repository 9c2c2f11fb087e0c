#!/usr/bin/env python3
"""Convert a pbrt `Shape "trianglemesh"` text block into the flat binary mesh the
scene files in scenes/ reference:  int32 nv, nt, hasN, hasUV; P[3nv] f32; N[3nv] f32;
uv[2nv] f32; idx[3nt] i32.

Used once, in the build container, on the head mesh the reference ships as a scene
ASSET (scenes/geometry/head.pbrt: 8 244 vertices, 15 864 triangles) -- the input the
reference's headline skin.pbrt render is quoted on.  The binary is
committed as scenes/head.mesh.bin so the GPU box (which has no /root/reference) can
run the same workload.  Values are parsed with float32(atof) exactly as pbrt's lexer
does (core/pbrtlex.ll NUMBER rule -> atof -> float).
"""
import re
import sys
import numpy as np


def block(text, name):
    m = re.search(r'"%s"\s*\[(.*?)\]' % re.escape(name), text, re.S)
    return None if m is None else m.group(1).split()


def main(src, dst):
    text = open(src).read()
    P = np.array([float(x) for x in block(text, "point P")], dtype=np.float64).astype(np.float32)
    idx = np.array([int(x) for x in block(text, "integer indices")], dtype=np.int32)
    N = block(text, "normal N")
    uv = block(text, "float uv") or block(text, "float st")
    nv, nt = P.size // 3, idx.size // 3
    with open(dst, "wb") as f:
        np.array([nv, nt, 1 if N else 0, 1 if uv else 0], dtype=np.int32).tofile(f)
        P.tofile(f)
        if N:
            np.array([float(x) for x in N], dtype=np.float64).astype(np.float32).tofile(f)
        if uv:
            np.array([float(x) for x in uv], dtype=np.float64).astype(np.float32).tofile(f)
        idx.tofile(f)
    print(f"{dst}: {nv} verts, {nt} tris, N={bool(N)}, uv={bool(uv)}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
