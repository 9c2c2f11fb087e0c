#!/usr/bin/env python3
"""Mo stage of ONE rank's share of the skin.scene frame (render tasks of rank r of w, as the multi-GPU path deals them) on one
GPU, for the packet sizes of k_mo_lambda: how much of a short launch is the tail of its heaviest packets."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package()
out = {}
for hits in (32, 16, 8, 0):
    os.environ["SKC_MO_PACKET_HITS"] = str(hits)
    dev = pkg.SkinDevice(0)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin.scene"))
    hs.upload(dev); hs.build_profile(dev)
    n_pts = dev.tessellate_resident(hs.minsampledistance, material_id=0)
    dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
    dev.build_octree()
    ref = None
    for world, rank in ((1, 0), (2, 0), (4, 1), (8, 0), (8, 5)):
        n = dev.frame_setup(hs.frame_desc(n_cores=1, rank=rank, world=world)); dev.frame_samples()
        ms = []
        for _ in range(3):
            dev.frame_film_clear(); L = dev.frame_render(0, n, hs.maxerror, hs.mix, want_L=(world == 8 and rank == 0)); ms.append(dev.stage_ms("mo"))
        out.setdefault("world %d rank %d (%d samples)" % (world, rank, n), {})["packet %d" % hits] = round(float(np.median(ms)), 3)
        if world == 8 and rank == 0:
            out.setdefault("_L", {})[hits] = L
    hs.close(); dev.close()
Ls = out.pop("_L")
out["rows identical across packet sizes"] = bool(all(np.array_equal(Ls[32].view(np.uint32), Ls[h].view(np.uint32)) for h in Ls))
print(json.dumps(out, indent=1))
