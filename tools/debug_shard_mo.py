"""Mo stage of one rank's share of the frame on a single GPU: python tools/debug_shard_mo.py WORLD (render tasks of rank 0 of WORLD).
Used to tune the warps-per-packet choice of short launches (SKC_MO_SPLIT=1|2|4 forces it)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package()
world = int(sys.argv[1])
dev = pkg.SkinDevice(0)
scene = pkg.HostScene(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes", "skin.scene"))
scene.upload(dev); scene.build_profile(dev)
n = dev.tessellate_resident(scene.minsampledistance, material_id=0)
dev.irradiance_resident(0, n, scene.irradiance_tasks(n, 1), scene.mix); dev.build_octree()
r = dev.frame_setup(scene.frame_desc(n_cores=1, rank=0, world=world))
dev.frame_samples(); dev.frame_film_clear()
ms = []
for _ in range(4):
    dev.frame_render(0, r, scene.maxerror, scene.mix); ms.append(dev.stage_ms("mo"))
print("world", world, "split", os.environ.get("SKC_MO_SPLIT", "auto"), "samples", r, "mo ms", round(float(np.median(ms)), 2))
