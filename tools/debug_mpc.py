"""Dump skc_profile_batch outputs for the known-answer cases (debug aid; run on the GPU box)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from __graft_entry__ import load_package
from golden_util import load_mpc_cases
pkg = load_package()
dev = pkg.SkinDevice(0)
out = {}
for k, (specs, opt, res, totals, ref) in enumerate(load_mpc_cases()):
    d2, R, T, tot = dev.profile_batch(specs, [len(specs)], [opt], resample=bool(res))
    out["R%d" % k] = R[0]; out["T%d" % k] = T[0]; out["tot%d" % k] = tot[0]
np.savez_compressed("gpurun_out/mpc_dbg.npz", **out)
