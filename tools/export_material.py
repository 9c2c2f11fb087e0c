#!/usr/bin/env python3
"""INTERIM: export the LayeredSkin material tables (Rd profile, rho_hd) of a scene with the
reference checker (oracle/_ref/ref_driver) into scenes/_generated/<scene>_tables.npz.

Used only until the device profile generator (skc_profile_batch) and the rho table generator
cover the material construction; bench.py prefers device-generated tables when available.
The output is git-ignored (it is derived data) but travels with gpurun."""
import os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
def main(scene):
    src = os.path.join(ROOT, "scenes", scene + ".scene")
    txt = open(src).read().replace("film 1280 720", "film 16 9")
    txt = "\n".join(l if not l.startswith("integrator") else "integrator maxdepth 5 maxerror 0.1 minsampledistance 0.2" for l in txt.splitlines())
    with tempfile.TemporaryDirectory() as d:
        sp = os.path.join(os.path.dirname(src), "_tmp_export.scene")
        open(sp, "w").write(txt)
        try:
            subprocess.run([os.path.join(ROOT, "oracle/_ref/ref_driver"), "scene", sp, d, "--ncores", str(os.cpu_count())], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        finally:
            os.remove(sp)
        f32 = lambda fn: np.fromfile(os.path.join(d, fn), np.float32)
        out = os.path.join(ROOT, "scenes", "_generated"); os.makedirs(out, exist_ok=True)
        np.savez(os.path.join(out, scene + "_tables.npz"), profile_hdr=f32("profile_hdr.bin").reshape(30, 4),
                 profile_rd=f32("profile_rd.bin").reshape(30, -1), rho_hd=f32("rho_hd.bin").reshape(-1, 30), skin_coeffs=f32("skin_coeffs.bin"))
        print("wrote", os.path.join(out, scene + "_tables.npz"))
if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "skin")
