#!/usr/bin/env python3
"""Per-wavelength windows of the compact Rd table for a scene's profile tables (host builder; tables come from the device)."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge

pkg = ge.load_package()
dev = pkg.SkinDevice(0)
hs = pkg.HostScene(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "scenes", "skin.scene"))
hs.upload(dev)
hdr, rd = hs.build_profile(dev)
out = {}
for tol in (1e-5, 5e-5):
    cp = pkg.skc.CompactProfile(rd, hdr[:, 1], tol=tol, head_cap=0)
    out[str(tol)] = dict(f_lo=cp.f_lo.tolist(), f_hi=cp.f_hi.tolist(), coverage=cp.coverage, rows=cp.rows)
    cp.close()
out["len"] = int(rd.shape[1]); out["rcp"] = hdr[:, 1].tolist()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "mo_windows.json"), "w"))
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "skin_tables.npz"), rd=rd, hdr=hdr)
hs.close(); dev.close()
