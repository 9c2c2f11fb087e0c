#!/usr/bin/env python3
"""A/B of the Mo kernel on one frame of a scene: exact tables only (k_mo_packet, SKC_MO_TABLE=exact) against the compact
shared-memory table with exact fallback (k_mo_lambda).  Prints the Mo stage times, the table statistics and the
per-sample difference of the radiance rows; writes gpurun_out/mo_table_ab.json."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def frame(pkg, scene_path, mode, reps=3, counters=True, env=None):
    os.environ["SKC_MO_TABLE"] = mode
    for k in ("SKC_MO_GROUP", "SKC_MO_WARPS", "SKC_MO_TOL", "SKC_MO_HEADCAP", "SKC_MO_WIDE"):
        os.environ.pop(k, None)
    os.environ.update(env or {})
    dev = pkg.SkinDevice(0)
    hs = pkg.HostScene(scene_path)
    try:
        hs.upload(dev)
        hs.build_profile(dev)
        n_pts = dev.tessellate_resident(hs.minsampledistance, material_id=0)
        dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
        dev.build_octree()
        n = dev.frame_setup(hs.frame_desc(n_cores=1))
        dev.frame_samples()
        ms = []
        L = None
        for _ in range(reps):
            dev.frame_film_clear()
            L = dev.frame_render(0, n, hs.maxerror, hs.mix, want_L=True)
            ms.append(dev.stage_ms("mo"))
        rgb = hs.film_to_rgb(dev.frame_film())
        st = {}
        if counters:
            c = dev.camera_pass_resident(0, n, hs.maxerror, hs.mix, counters=True)
            st = dev.mo_table_stats()
            st["counters7"] = [int(x) for x in c]
            st["group_hist"] = dev.mo_group_histogram()
        return dict(ms=ms, L=L, rgb=rgb, stats=st, n_pts=n_pts, n=n)
    finally:
        hs.close()
        dev.close()


def main():
    pkg = ge.load_package()
    scene = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "scenes", "skin.scene")
    a = frame(pkg, scene, "exact")
    b = frame(pkg, scene, "compact")
    La, Lb = a["L"].astype(np.float64), b["L"].astype(np.float64)
    scale = np.abs(La).max(axis=1)
    lit = scale > 0
    err = np.abs(La - Lb).max(axis=1)[lit] / scale[lit]
    rmse = float(np.sqrt(np.mean((a["rgb"] - b["rgb"]) ** 2)) / np.sqrt(np.mean(a["rgb"] ** 2)))
    out = {"scene": os.path.basename(scene), "points": a["n_pts"], "samples": a["n"],
           "mo_ms_exact": a["ms"], "mo_ms_compact": b["ms"],
           "L_row_rel_err_max": float(err.max()), "L_row_rel_err_p999": float(np.quantile(err, 0.999)), "L_row_rel_err_mean": float(err.mean()),
           "image_rmse_rel": rmse, "stats_exact": a["stats"], "stats_compact": b["stats"]}
    out["variants"] = {}
    variants = [("tol3e-5", {"SKC_MO_TOL": "3e-5"}), ("tol5e-5", {"SKC_MO_TOL": "5e-5"}), ("tol8e-5", {"SKC_MO_TOL": "8e-5"})]
    for name, env in variants:
        if os.environ.get("MO_AB_VARIANTS", "1") == "0":
            break
        c = frame(pkg, scene, "compact", counters=True, env=env)
        e2 = np.abs(La - c["L"].astype(np.float64)).max(axis=1)[lit] / scale[lit]
        out["variants"][name] = {"mo_ms": c["ms"], "L_row_rel_err_max": float(e2.max()), "stats": c["stats"],
                                 "identical_to_default": bool(np.array_equal(c["L"].view(np.uint32), b["L"].view(np.uint32)))}
    print(json.dumps(out, indent=1))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "mo_table_ab.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
