#!/usr/bin/env python3
"""BASELINE.json configs[3] on the B200: the synthetic Mo sweep and the BVH microbenchmark.

  Mo sweep   unit icosphere (level 8: 1.3 M triangles), LayeredSkin defaults, N_pts in {1 M, 4 M, 16 M, 64 M} surface points from
             the device tessellator (minsampledistance chosen so the count lands nearest the target: the per-edge factors are
             integers, so the counts are quantised), the analytic irradiance field of tools/synthetic.py, octree build, then Mo for
             a coherent hit set (2048^2 camera x 1 / 4 spp through the frame path) and an incoherent one (random surface
             points), maxError in {0.05, 0.1, 0.2}
  BVH        (i) the irradiance pass's own shadow rays (coherent any-hit); (ii) 16 M rays of the AggregateTest distribution,
             closest hit and any hit; on the head mesh (15 864 triangles) and on icospheres of 1.3 M / 5.2 M (/ 21 M) triangles

Device times are the library's stage timers (CUDA events on the launching stream, inputs resident in HBM); algorithmic bytes
follow SURVEY.md 8d: BVH 40 B/ray + 32 B/slab test + 48 B/primitive test + 16 B/closest hit or 1 B/any hit; Mo as in bench.py.
One JSON document on stdout (-> profiles/)."""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import synthetic as syn  # noqa: E402


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def pick_min_dist(scene, target):
    """minsampledistance whose tessellation count is nearest the target (log scale)."""
    lo, hi = 1e-4, 0.1
    best = None
    for _ in range(40):
        mid = (lo * hi) ** 0.5
        n = scene.count_points(mid)
        if best is None or abs(np.log(n / target)) < abs(np.log(best[1] / target)):
            best = (mid, n)
        if n > target:
            lo = mid
        else:
            hi = mid
    return best


def timed(dev, fn, stage, reps=3):
    fn()
    ms = []
    for _ in range(reps):
        fn(); ms.append(dev.stage_ms(stage))
    return float(np.median(ms))


REF_DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")


def ref_trace(scene_path, rays, any_hit, tmp):
    """The reference's own Scene::Intersect / IntersectP on the same rays, all host cores (oracle/_ref/ref_driver trace)."""
    import subprocess
    rp = os.path.join(tmp, "rays.bin")
    np.ascontiguousarray(rays, np.float32).tofile(rp)
    out = subprocess.run([REF_DRIVER, "trace", scene_path, rp, "--reps", "2"] + (["--any"] if any_hit else []), capture_output=True, text=True).stdout
    cores = int([l for l in out.splitlines() if l.startswith("cores")][0].split()[1])
    rows = [l.split() for l in out.splitlines() if l.startswith("rays")]
    sec = min(float(r[5]) for r in rows)
    return {"cores": cores, "rays": int(rows[0][1]), "hits": int(rows[0][3]), "seconds": sec, "Mrays_per_s": int(rows[0][1]) / sec / 1e6}


def ref_mo(scene_path):
    """The reference's Preprocess + a camera pass that evaluates only the subsurface term (ref_driver bench --sss-only), all host cores."""
    import subprocess
    out = subprocess.run([REF_DRIVER, "bench", scene_path, "--sss-only", "--steps", "1", "--warmup", "0"], capture_output=True, text=True).stdout
    cores = int([l for l in out.splitlines() if l.startswith("cores")][0].split()[1])
    w = [l for l in out.splitlines() if l.startswith("step")][0].split()
    return {"cores": cores, "points": int(w[3]), "samples": int(w[5]), "preprocess_s": float(w[7]), "camera_pass_sss_only_s": float(w[9]),
            "Mhits_per_s_incl_camera_rays": int(w[5]) / float(w[9]) / 1e6}


def bvh_bench(pkg, dev, scene, name, n_rays, reps, ref_scene=None, tmp=None):
    PEAK = peak()
    out = {"mesh": name, "triangles": scene.n_tris, "bvh_nodes": scene.n_nodes, "bvh_bytes": int(scene.n_nodes * 32 + scene.n_prims * 48)}
    bmin = scene.P.min(axis=0); bmax = scene.P.max(axis=0)          # the mesh's bound (the emitters sit far outside it)
    for dist, rays in (("aggregate", syn.aggregate_test_rays(bmin, bmax, n_rays)), ("through_mesh", syn.through_mesh_rays(scene.P, n_rays))):
        dev.set_camera_rays(rays)
        for mode, any_hit, per_ray in (("closest", False, 40 + 16), ("any", True, 40 + 1)):
            hits, c = dev.trace_resident(0, n_rays, any_hit=any_hit, counters=True, want_out=True)
            frac_hit = float(np.mean(hits != 0)) if any_hit else float(np.mean(hits["hit"] != 0))
            ms = timed(dev, lambda: dev.trace_resident(0, n_rays, any_hit=any_hit), "any" if any_hit else "closest", reps)
            byts = n_rays * per_ray + 32 * int(c[0]) + 48 * int(c[1])
            out[dist + "_" + mode] = {"rays": n_rays, "hit_fraction": frac_hit, "ms": ms, "Mrays_per_s": n_rays / ms / 1e3, "slab_tests_per_ray": int(c[0]) / n_rays,
                                      "prim_tests_per_ray": int(c[1]) / n_rays, "algorithmic_GBps": byts / ms / 1e6, "frac_of_hbm_peak": byts / ms / 1e6 / PEAK}
            if ref_scene:       # the same rays (the first 4 M of them) through the reference on the host cores; the hit counts must agree
                m = min(n_rays, 4 << 20)
                cpu = ref_trace(ref_scene, rays[:m], any_hit, tmp)
                got = int(np.count_nonzero(hits[:m] != 0)) if any_hit else int(np.count_nonzero(hits["hit"][:m] != 0))
                cpu["hit_count_equals_device"] = bool(cpu["hits"] == got)
                out[dist + "_" + mode]["cpu_reference"] = cpu
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--targets", default="1e6,4e6,16e6")
    ap.add_argument("--level", type=int, default=8)
    ap.add_argument("--bvh-levels", default="8,9")
    ap.add_argument("--rays", type=int, default=16 << 20)
    ap.add_argument("--hits", type=int, default=4 << 20)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--reference", action="store_true", help="time the reference (oracle/_ref/ref_driver) beside the device: the first Mo sweep size and every BVH case")
    args = ap.parse_args()
    if args.reference and not os.path.exists(REF_DRIVER):
        sys.exit("--reference: oracle/_ref/ref_driver is not built")
    if args.quick:
        args.targets, args.level, args.bvh_levels, args.rays, args.hits = "2e5", 5, "5", 1 << 18, 1 << 16
    import __graft_entry__ as ge
    pkg = ge.load_package()
    dev = pkg.SkinDevice(0)
    PEAK = peak()
    tmp = tempfile.mkdtemp(prefix="skin_synth_")
    result = {"hbm_peak_GBps": PEAK, "mo_sweep": [], "bvh": []}

    # ---------------------------------------------------------------- Mo sweep
    t0 = time.time()
    film = 256 if args.quick else 2048
    path = syn.write_scene(tmp, args.level, film=film, spp=1)
    scene = pkg.HostScene(path)
    result["icosphere"] = {"level": args.level, "triangles": scene.n_tris, "bvh_nodes": scene.n_nodes, "host_bvh_build_s": round(time.time() - t0, 2)}
    scene.upload(dev)
    scene.build_profile(dev)
    for target in [float(x) for x in args.targets.split(",")]:
        md, _ = pick_min_dist(scene, target)
        n_pts = dev.tessellate_resident(md, False, material_id=2)
        sp = dev.get_surface_points(0, n_pts)
        ip = syn.synthetic_irradiance(sp, pkg.DT_IRR_POINT)
        dev.set_irradiance_points(ip)
        del ip
        ms_tree = timed(dev, dev.build_octree, "octree", args.reps)
        n_nodes = dev.build_octree()[0]
        row = {"target_points": int(target), "points": int(n_pts), "minsampledistance": md, "octree_nodes": int(n_nodes), "octree_ms": ms_tree,
               "octree_Mpoints_per_s": n_pts / ms_tree / 1e3, "runs": []}
        hit_sets = []
        for spp in (1, 4):
            fd = scene.frame_desc(n_cores=1)
            fd[8:12].view(np.int32)[0] = spp
            hit_sets.append(("camera %dx%dx%d (coherent, frame path, Z-order packets)" % (film, film, spp), fd))
        hit_sets.append(("random surface points (incoherent)", syn.random_hit_points(sp, args.hits)))
        del sp
        for label, hs in hit_sets:
            for max_err in (0.05, 0.1, 0.2):
                if isinstance(hs, np.ndarray) and hs.ndim == 2:
                    n_hit = len(hs)
                    dev.set_camera_rays(np.zeros((1, 8), np.float32))
                    ms = timed(dev, lambda: dev.mo(hs, max_err), "mo", args.reps)
                    cnt = None
                else:
                    n_hit = dev.frame_setup(hs)
                    dev.frame_samples()
                    dev.frame_film_clear()
                    ms = timed(dev, lambda: dev.frame_render(0, n_hit, max_err, 0.5), "mo", args.reps)
                    cnt = dev.camera_pass_resident(0, n_hit, max_err, 0.5, counters=True)
                run = {"hits": label, "n_hit": int(n_hit), "maxError": max_err, "mo_ms": ms, "Mhits_per_s": n_hit / ms / 1e3}
                if cnt is not None:
                    byts = n_hit * (12 + 128) + 48 * int(cnt[2]) + 16 * int(cnt[3]) + 368 * (int(cnt[4]) + int(cnt[5]))
                    run.update({"nodes_per_hit": int(cnt[2]) / n_hit, "contributions_per_hit": (int(cnt[4]) + int(cnt[5])) / n_hit,
                                "algorithmic_GBps": byts / ms / 1e6, "frac_of_hbm_peak": byts / ms / 1e6 / PEAK,
                                "table_gathers_per_s_G": int(cnt[6]) / ms / 1e6})
                row["runs"].append(run)
        if args.reference and not result["mo_sweep"]:
            # the same icosphere, point set and camera through the reference; its irradiance comes from the scene's two lights
            # instead of the analytic field (the octree walk depends on positions and areas only, not on E)
            rp = syn.write_scene(os.path.join(tmp, "ref"), args.level, min_dist=md, film=film, spp=1)
            row["cpu_reference_maxError_0.1_spp1"] = ref_mo(rp)
        result["mo_sweep"].append(row)
        print("mo sweep: %d points done (%.0f s)" % (n_pts, time.time() - t0), file=sys.stderr)

    # ---------------------------------------------------------------- BVH: the irradiance pass's shadow rays
    md, _ = pick_min_dist(scene, 4e6 if not args.quick else 2e5)
    n_pts = dev.tessellate_resident(md, False, material_id=2)
    n_tasks = scene.irradiance_tasks(n_pts, 1)
    c = dev.irradiance_resident(0, n_pts, n_tasks, 0.5, counters=True)
    ms = timed(dev, lambda: dev.irradiance_resident(0, n_pts, n_tasks, 0.5), "irradiance", args.reps)
    byts = n_pts * (44 + 144 + 24) + int(c[2]) * 41 + 32 * int(c[0]) + 48 * int(c[1])
    result["shadow_rays_icosphere"] = {"points": int(n_pts), "shadow_rays": int(c[2]), "irradiance_pass_ms": ms, "Mrays_per_s_whole_pass": int(c[2]) / ms / 1e3,
                                       "slab_tests_per_ray": int(c[0]) / max(1, int(c[2])), "prim_tests_per_ray": int(c[1]) / max(1, int(c[2])),
                                       "algorithmic_GBps": byts / ms / 1e6}
    scene.close()

    # ---------------------------------------------------------------- BVH: AggregateTest rays on meshes of growing size
    head = pkg.HostScene(os.path.join(ROOT, "scenes", "skin.scene"))
    head.upload(dev)
    head_ref = None
    if args.reference:      # skin.scene with a short profile: only the mesh matters to the ray loop
        head_ref = os.path.join(tmp, "head_ref.scene")
        src = open(os.path.join(ROOT, "scenes", "skin.scene")).read().replace(" layers ", " desiredlength 32 layers ").replace("\nmesh ", "\nmesh " + os.path.join(ROOT, "scenes") + "/")
        open(head_ref, "w").write(src)
    result["bvh"].append(bvh_bench(pkg, dev, head, "head (scenes/skin.scene)", args.rays, args.reps, head_ref, tmp))
    head.close()
    for lvl in [int(x) for x in args.bvh_levels.split(",")]:
        t1 = time.time()
        sp_path = syn.write_scene(tmp, lvl, film=64, desired_length=32 if args.reference else 512)
        sc = pkg.HostScene(sp_path)
        build_s = time.time() - t1
        sc.upload(dev)
        r = bvh_bench(pkg, dev, sc, "icosphere level %d" % lvl, args.rays, args.reps, sp_path if args.reference else None, tmp)
        r["host_bvh_build_s"] = round(build_s, 2)
        result["bvh"].append(r)
        sc.close()
        print("bvh: level %d done (%.0f s)" % (lvl, time.time() - t0), file=sys.stderr)
    print(json.dumps(result, indent=1))
    dev.close()


if __name__ == "__main__":
    main()
