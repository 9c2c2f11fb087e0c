#!/usr/bin/env python3
"""Developer aid: Mo time of each rank's share of the frame (round-robin render tasks), on one GPU."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g
pkg = g.load_package()
dev = pkg.SkinDevice(0)
hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin.scene"))
hs.upload(dev)
hs.build_profile(dev)
n_pts = dev.tessellate_resident(hs.minsampledistance, False, material_id=0)
dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
dev.build_octree()
for world in (1, 2, 4, 8):
    for rank in range(world):
        n = dev.frame_setup(hs.frame_desc(1, rank, world))
        for _ in range(2):
            dev.frame_samples()
            dev.frame_render(0, n, hs.maxerror, hs.mix)
        dev.synchronize()
        print("world %d rank %d: samples %d  mo %.2f ms  samples-stage %.2f  direct %.2f  film %.2f" % (world, rank, n, dev.stage_ms("mo"), dev.stage_ms("samples"), dev.stage_ms("direct"), dev.stage_ms("film")))
