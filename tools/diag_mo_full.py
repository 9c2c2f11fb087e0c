"""Diagnosis: device octree / Mo at full skin.scene size against tests/golden/skin_full.npz, for each builder and Mo kernel."""
import os, sys, json, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
import golden_util as gu
pkg = ge.load_package()
z = np.load(os.path.join(ROOT, "tests", "golden", "skin_full.npz"))
hits = z["hits"].view(gu.DT_HIT).reshape(-1); s = np.nonzero(hits["hasBssrdf"] == 1)[0]
for env in ({},):
    for k in ("SKC_OCTREE_LEVELWISE", "SKC_MO_TABLE", "SKC_MO_KERNEL"):
        os.environ.pop(k, None)
    os.environ.update(env)
    dev = pkg.SkinDevice(0); hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin.scene"))
    hs.upload(dev); hs.build_profile(dev, rho_hd=(z['rho_hd'] if env.get('REFRHO') else None))
    n = dev.tessellate_resident(hs.minsampledistance, material_id=0)
    dev.irradiance_resident(0, n, hs.irradiance_tasks(n, 1), hs.mix)
    if env.get("PATCH"):
        for i, row in zip(z["irradiance_flip_idx"], z["irradiance_flip_rows"]):
            dev.set_irradiance_rows(int(i), row[6:36])
    pa, E = dev.irradiance_rows(0, len(z["irradiance_points"]), int(z["irradiance_stride"]))
    print("  sampled irradiance rows bit-exact:", float(np.mean((E.view(np.uint32) == z["irradiance_points"][:, 6:36].view(np.uint32)).all(axis=1))))
    nn, root, b = dev.build_octree()
    mo = dev.mo(hits["p"][s], hs.maxerror)
    e = gu.row_rel_err(mo, z["mo"][s])
    bad = np.nonzero(e > 1e-4)[0]
    print(env, "nodes", nn, "(oracle 8125227) root maxabsdiff", float(np.abs(root - z["octree_root"]).max()), "root rel", float((np.abs(root - z["octree_root"]) / np.abs(z["octree_root"])).max()),
          "Mo max", float(e.max()), "bad rows", [(int(i), float(e[i])) for i in bad][:20])
    hs.close(); dev.close()
