#!/usr/bin/env python3
"""Developer aid: where does the device LDSampler stream differ from the golden export?"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g
pkg = g.load_package()
z = np.load(os.path.join(ROOT, "tests", "golden", "skin_ld4.npz"))
dev = pkg.SkinDevice(0)
hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_ld4.scene"))
hs.upload(dev)
n = dev.frame_setup(hs.frame_desc(1))
dev.frame_samples()
xy, arr, rays = dev.frame_get_samples(0, n)
bx = (xy.view(np.uint32) != z["camera_xy"].view(np.uint32)).any(axis=1)
print("xy mismatches", bx.sum(), "of", n, "first", np.flatnonzero(bx)[:10])
for i in np.flatnonzero(bx)[:5]:
    print(i, xy[i], z["camera_xy"][i])
ba = (arr.view(np.uint32) != z["camera_samples"].view(np.uint32))
print("array mismatches", ba.sum(), "rows", ba.any(axis=1).sum(), "cols", np.flatnonzero(ba.any(axis=0)))
i = np.flatnonzero(ba.any(axis=1))[:3]
for k in i:
    print(k, arr[k][:8], z["camera_samples"][k][:8])
br = (rays[:, :6].view(np.uint32) != z["camera_rays"][:, :6].view(np.uint32)).any(axis=1)
print("ray mismatches", br.sum())
# ---- radiance
hs.build_profile(dev)
n_pts = dev.tessellate_resident(hs.minsampledistance, False, material_id=0)
dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
dev.build_octree()
Lr = dev.frame_render(0, n, hs.maxerror, hs.mix, want_L=True)
order = dev.frame_order(0, n)
L = np.empty_like(Lr); L[order] = Lr
Lsr = np.empty((n, 30), np.float32)
dev.camera_pass_resident(0, n, hs.maxerror, hs.mix, out=Lsr)
dev.synchronize()
Ls = np.empty_like(Lsr); Ls[order] = Lsr
dev.synchronize()
ref, ref_s = z["L"], z["L_sss"]
d_ref, d_got = ref - ref_s, L - Ls
sc = np.abs(ref).max(axis=1) + 1e-30
lit = ref.max(axis=1) > 0
es = np.abs(Ls - ref_s).max(axis=1) / sc
ed = np.abs(d_got - d_ref).max(axis=1) / sc
print("sss term: within 1e-4 %.4f max %.3g" % (np.mean(es[lit] < 1e-4), es[lit].max()))
print("direct term: within 1e-4 %.4f within 1e-5 %.4f max %.3g" % (np.mean(ed[lit] < 1e-4), np.mean(ed[lit] < 1e-5), ed[lit].max()))
worst = np.argsort(-ed * lit)[:8]
for i in worst:
    print(i, "direct ref %.6g got %.6g  total %.6g  rel %.3g" % (d_ref[i, 10], d_got[i, 10], ref[i, 10], ed[i]))
rat = d_got[lit, 10] / np.where(d_ref[lit, 10] != 0, d_ref[lit, 10], 1)
print("ratio got/ref direct: median %.6f  p5 %.6f p95 %.6f" % (np.median(rat), np.percentile(rat, 5), np.percentile(rat, 95)))
print("direct share of total: median %.4f" % np.median(d_ref[lit, 10] / ref[lit, 10]))
