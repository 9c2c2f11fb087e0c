"""Wall-clock split of the host-buffer (e2e) step: irradiance / octree / camera pass (debug aid, GPU box)."""
import os, sys, time
import numpy as np
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package
pkg = load_package()
dev = pkg.SkinDevice(0)
scene = pkg.HostScene(os.path.join(ROOT, "scenes", "skin.scene"))
scene.upload(dev)
scene.build_profile(dev, None, upload=True)
pts = scene.tessellate(); rays = scene.bench_rays()
n_tasks = scene.irradiance_tasks(len(pts), 1)
h_pts = dev.pinned((len(pts),), pkg.DT_SURFACE_POINT); h_pts[:] = pts
h_rays = dev.pinned((len(rays), 8), np.float32); h_rays[:] = rays
h_L = dev.pinned((len(rays), 30), np.float32)
for it in range(4):
    t0 = time.perf_counter()
    dev.irradiance(h_pts, n_tasks, scene.mix, first=0, count=len(pts), want_out=False)
    t1 = time.perf_counter()
    dev.build_octree()
    t2 = time.perf_counter()
    dev._ck(dev.lib.skc_camera_pass(dev.ctx, h_rays.ctypes.data, len(rays), scene.maxerror, scene.mix, None, h_L.ctypes.data))
    t3 = time.perf_counter()
    print("irradiance %.1f ms (stage %.1f)  octree %.1f  camera %.1f (mo stage %.1f)" % ((t1 - t0) * 1e3, dev.stage_ms("irradiance"), (t2 - t1) * 1e3, (t3 - t2) * 1e3, dev.stage_ms("mo")))
