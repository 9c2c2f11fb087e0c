#!/usr/bin/env python3
"""One frame of a scene up to and including the Mo pass (for profiling k_mo_lambda under ncu with the environment's knobs)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
dev = pkg.SkinDevice(0)
hs = pkg.HostScene(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "scenes", "skin.scene"))
hs.upload(dev); hs.build_profile(dev)
n_pts = dev.tessellate_resident(hs.minsampledistance, material_id=0)
dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
dev.build_octree()
n = dev.frame_setup(hs.frame_desc(n_cores=1)); dev.frame_samples()
for _ in range(int(os.environ.get("MO_REPS", "2"))):
    dev.frame_film_clear(); dev.frame_render(0, n, hs.maxerror, hs.mix)
    print("mo ms", dev.stage_ms("mo"))
hs.close(); dev.close()
