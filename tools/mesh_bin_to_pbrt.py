#!/usr/bin/env python3
"""Write the flat binary mesh of scenes/*.mesh.bin (see tools/pbrt_mesh_to_bin.py) back out as a pbrt
`Shape "trianglemesh"` text block, so that the `.pbrt` scene files in scenes/ can `Include "geometry/head.pbrt"` exactly as the
reference's skin scenes do (results/PigmentManips/Eum0.1_Deoxy0.5.pbrt:47).  Numbers are printed with 9 significant digits:
(float)atof() of each returns the stored float32 bit for bit.  The output is a build product (scenes/geometry/ is git-ignored)."""
import os
import sys
import numpy as np


def convert(src, dst, extra=""):
    """extra: further parameters of the Shape directive, e.g. '"texture alpha" "alphamap"'."""
    raw = open(src, "rb").read()
    nv, nt, has_n, has_uv = np.frombuffer(raw, np.int32, 4)
    off = 16
    P = np.frombuffer(raw, np.float32, 3 * nv, off).reshape(-1, 3); off += 12 * nv
    N = uv = None
    if has_n:
        N = np.frombuffer(raw, np.float32, 3 * nv, off).reshape(-1, 3); off += 12 * nv
    if has_uv:
        uv = np.frombuffer(raw, np.float32, 2 * nv, off).reshape(-1, 2); off += 8 * nv
    idx = np.frombuffer(raw, np.int32, 3 * nt, off).reshape(-1, 3)
    os.makedirs(os.path.dirname(dst) or ".", exist_ok=True)

    def rows(a):
        return "\n".join("      " + " ".join("%.9g" % float(v) for v in r) for r in a)
    with open(dst, "w") as f:
        f.write('AttributeBegin\n  Shape "trianglemesh"\n    "point P" [\n%s\n    ]\n' % rows(P))
        if N is not None:
            f.write('    "normal N" [\n%s\n    ]\n' % rows(N))
        if uv is not None:
            f.write('    "float uv" [\n%s\n    ]\n' % rows(uv))
        f.write('    "integer indices" [\n%s\n    ]\n' % "\n".join("      %d %d %d" % tuple(r) for r in idx))
        if extra:
            f.write("    %s\n" % extra)
        f.write("AttributeEnd\n")
    return int(nv), int(nt)


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    src = sys.argv[1] if len(sys.argv) > 1 else os.path.join(here, "..", "scenes", "head.mesh.bin")
    dst = sys.argv[2] if len(sys.argv) > 2 else os.path.join(here, "..", "scenes", "geometry", "head.pbrt")
    print("%s: %d vertices, %d triangles" % ((dst,) + convert(src, dst)))
