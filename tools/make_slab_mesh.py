#!/usr/bin/env python3
"""The mesh of BASELINE.json configs[0]'s stand-in scene (scenes/tissue.pbrt is absent from the fork):
a 1 x 1 x 0.1 slab given as two quads (top z = 0.1 facing +z, bottom z = 0 facing -z), each tessellated to an n x n grid
(n = 70 -> 19 600 triangles), no vertex normals, no uvs.  Written in the flat binary layout of tools/pbrt_mesh_to_bin.py
and as a pbrt `Shape "trianglemesh"` block.  Deterministic; the outputs are build products (scenes/geometry/ is git-ignored)."""
import os
import sys
import numpy as np


def slab(n=70):
    g = np.arange(n + 1, dtype=np.float64) / n
    x, y = np.meshgrid(g, g, indexing="xy")
    top = np.stack([x.ravel(), y.ravel(), np.full(x.size, 0.1)], 1)
    bot = np.stack([x.ravel(), y.ravel(), np.zeros(x.size)], 1)
    P = np.concatenate([top, bot]).astype(np.float32)
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="xy")
    v00 = (j * (n + 1) + i).ravel(); v10 = v00 + 1; v01 = v00 + n + 1; v11 = v01 + 1
    t_top = np.stack([np.stack([v00, v10, v11], 1), np.stack([v00, v11, v01], 1)], 1).reshape(-1, 3)      # counter-clockwise from +z
    t_bot = t_top[:, ::-1] + (n + 1) ** 2
    return P, np.concatenate([t_top, t_bot]).astype(np.int32)


def write(dst_bin, dst_pbrt, n=70):
    P, idx = slab(n)
    for d in (dst_bin, dst_pbrt):
        os.makedirs(os.path.dirname(d) or ".", exist_ok=True)
    with open(dst_bin, "wb") as f:
        np.array([len(P), len(idx), 0, 0], np.int32).tofile(f); P.tofile(f); idx.tofile(f)
    with open(dst_pbrt, "w") as f:
        f.write('Shape "trianglemesh"\n  "point P" [\n%s\n  ]\n  "integer indices" [\n%s\n  ]\n' % (
            "\n".join("    %.9g %.9g %.9g" % tuple(float(v) for v in p) for p in P), "\n".join("    %d %d %d" % tuple(t) for t in idx)))
    return len(P), len(idx)


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    g = os.path.join(here, "..", "scenes", "geometry")
    print("slab: %d vertices, %d triangles" % write(os.path.join(g, "slab.mesh.bin"), os.path.join(g, "slab.pbrt")))
