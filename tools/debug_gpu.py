import sys, os
import numpy as np
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import __graft_entry__ as ge, golden_util as gu
pkg = ge.load_package()
g = gu.load_scene_golden('skin_tiny')
dev = pkg.SkinDevice(0)
gu.setup_device(dev, g)
hits = dev.trace_closest(g.camera_rays)
ref = g.hits
print('hit mismatches', (hits['hit'] != ref['hit']).sum(), 'of', len(ref), 'ref hits', ref['hit'].sum(), 'gpu hits', hits['hit'].sum())
m = (hits['hit'] == 1) & (ref['hit'] == 1)
print('prim mism', (hits['prim'][m] != ref['prim'][m]).sum(), 't mism', (hits['t'][m].view(np.uint32) != ref['t'][m].view(np.uint32)).sum())
bad = np.nonzero(hits['hit'] != ref['hit'])[0][:5]
for i in bad: print(i, g.camera_rays[i], ref[i], hits[i])
