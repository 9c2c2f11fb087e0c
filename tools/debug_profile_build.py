"""Build the skin.scene material tables once (debug / ncu target)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from __graft_entry__ import load_package
pkg = load_package(); dev = pkg.SkinDevice(0); sc = pkg.HostScene(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "scenes", "skin.scene"))
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 1):
    sc.build_profile(dev, None, upload=True); print("profile_ms", dev.stage_ms("profile"), "rho_ms", dev.stage_ms("rho"))
