"""Shared helpers for the parity tests: golden-fixture loading and device set-up."""
import os
import types

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

DT_HIT = np.dtype([("hit", "<i4"), ("prim", "<i4"), ("t", "<f4"), ("rayEpsilon", "<f4"), ("p", "<f4", 3), ("n", "<f4", 3), ("wo", "<f4", 3), ("hasBssrdf", "<i4")])
DT_SP = np.dtype([("p", "<f4", 3), ("n", "<f4", 3), ("u", "<f4"), ("v", "<f4"), ("materialId", "<u4"), ("area", "<f4"), ("rayEpsilon", "<f4")])
DT_IP = np.dtype([("p", "<f4", 3), ("n", "<f4", 3), ("E", "<f4", 30), ("area", "<f4"), ("rayEpsilon", "<f4")])
DT_NODE = np.dtype([("pmin", "<f4", 3), ("pmax", "<f4", 3), ("offset", "<u4"), ("nPrimitives", "u1"), ("axis", "u1"), ("pad", "u1", 2)])


def load_scene_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    g = types.SimpleNamespace()
    g.meta = dict(zip([str(k) for k in z["meta_keys"]], z["meta_vals"]))
    g.bvh_nodes = z["bvh_nodes"].view(DT_NODE)
    g.bvh_prims = z["bvh_prims"]
    g.mesh_P, g.mesh_N, g.mesh_uv, g.mesh_idx, g.mesh_o2w = z["mesh_P"], z["mesh_N"], z["mesh_uv"], z["mesh_idx"], z["mesh_o2w"]
    g.lights = z["lights"]
    g.surface_points = z["surface_points"].view(DT_SP)
    g.irradiance_points = np.ascontiguousarray(z["irradiance_points"]).view(DT_IP).reshape(-1)
    g.octree_bounds, g.octree_root = z["octree_bounds"], z["octree_root"]
    g.profile_hdr, g.profile_rd, g.rho_hd = z["profile_hdr"], z["profile_rd"], z["rho_hd"]
    g.skin_coeffs, g.cie = z["skin_coeffs"], z["cie_xyz30"]
    g.camera_rays, g.camera_xy = z["camera_rays"], z["camera_xy"]
    g.hits = z["hits"].view(DT_HIT)
    g.mo, g.L_sss, g.L = z["mo"], z["L_sss"], z["L"]
    g.max_error = float(g.meta["maxError"]); g.mix = float(g.meta["mix"])
    g.n_tasks = int(g.meta["irradiance_tasks"])
    return g


def setup_device(dev, g, tables=True):
    """Upload the exported flat BVH, lights and material tables of a scene golden."""
    dev.upload_bvh(g.bvh_nodes, g.bvh_prims, g.mesh_P, g.mesh_N, g.mesh_uv, g.mesh_idx, g.mesh_o2w,
                   reverse=int(g.meta["mesh0_reverse"]), swaps=int(g.meta["mesh0_swaps"]), spheres=g.lights)
    dev.upload_lights(g.lights)
    if tables:
        dev.upload_profile(g.profile_rd, g.profile_hdr[:, 1], g.rho_hd[:1025])


def rel_err(a, b, floor=0.0):
    """|a-b| / max(|b|, floor) elementwise (b = reference)."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), floor if floor > 0 else np.finfo(np.float64).tiny)


def row_rel_err(a, b):
    """Per-row relative error: max_c |a-b| / max_c |b| (spectra compared against their own scale)."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    scale = np.abs(b).max(axis=1)
    err = np.abs(a - b).max(axis=1)
    return np.where(scale > 0, err / np.where(scale > 0, scale, 1), err)


def smoke_check(pkg):
    g = load_scene_golden("skin_tiny")
    dev = pkg.SkinDevice(0)
    try:
        setup_device(dev, g)
        irr = dev.irradiance(g.surface_points, g.n_tasks, g.mix)
        e = row_rel_err(irr["E"], g.irradiance_points["E"])
        assert e.max() < 1e-6, "irradiance parity: max rel %.3g" % e.max()
        dev.build_octree()
        L = dev.camera_pass(g.camera_rays, g.max_error, g.mix)
        e = row_rel_err(L, g.L_sss)
        assert e.max() < 1e-5, "camera pass parity: max rel %.3g" % e.max()
        assert dev.launch_count() > 0
        # and against the C restatement on a few hits (oracle/skin_oracle.c)
        import oracle_api as orc
        dev.set_irradiance_points(g.irradiance_points)
        dev.build_octree()
        s = np.nonzero(g.hits["hasBssrdf"] == 1)[0][:64]
        tree = orc.Octree(g.irradiance_points, g.cie[1])
        ref = tree.mo(g.hits["p"][s], g.profile_rd, g.profile_hdr[:, 1], g.max_error)
        e = row_rel_err(dev.mo(g.hits["p"][s], g.max_error), ref)
        assert e.max() < 1e-5, "Mo vs oracle: %.3g" % e.max()
    finally:
        dev.close()
    # one whole frame from the scene file (device tessellation, sampler stream, full Li, film) against the reference's image
    z = np.load(os.path.join(GOLDEN, "skin_ld4.npz"))
    dev = pkg.SkinDevice(0)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_ld4.scene"))
    try:
        hs.upload(dev)
        hs.build_profile(dev)
        n_pts = dev.tessellate_resident(hs.minsampledistance, False, material_id=0)
        dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
        dev.build_octree()
        n = dev.frame_setup(hs.frame_desc(n_cores=1))
        dev.frame_samples()
        dev.frame_render(0, n, hs.maxerror, hs.mix)
        rgb = hs.film_to_rgb(dev.frame_film())
        ref = z["image"]
        rmse = np.sqrt(np.mean((rgb - ref) ** 2)) / np.sqrt(np.mean(ref ** 2))
        assert rmse < 1e-3, "frame RMSE %.3g relative" % rmse
    finally:
        hs.close()
        dev.close()


def load_mpc_cases():
    """tests/golden/mpc_kat.npz -> list of (specs, opts, resample, totals, d2_R_T); the step size follows
    core/multipole.cpp:262-268 (12 * mean mfp / desiredLength) in fp32, as the generating driver did."""
    z = np.load(os.path.join(GOLDEN, "mpc_kat.npz"))
    out = []
    i = 0
    while "case%d_spec" % i in z.files:
        s = z["case%d_spec" % i]
        nl, dl, lerp, res, step = int(s[0]), int(s[1]), int(s[2]), int(s[3]), np.float32(s[4])
        layers = s[5:].reshape(nl, 4).astype(np.float32)
        if not step > 0:
            tot = np.float32(0)
            for l in layers:
                tot = np.float32(tot + np.float32(1) / np.float32(l[2] + l[3]))
            mfp = np.float32(tot / np.float32(nl))
            step = np.float32(np.float32(np.float32(12) * mfp) / np.float32(dl))
        specs = [(l[0], l[1], l[2], l[3], 0.0) for l in layers]
        out.append((specs, (step, dl, lerp), res, z["case%d_totals" % i], z["case%d_d2_R_T" % i]))
        i += 1
    return out
