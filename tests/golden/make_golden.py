#!/usr/bin/env python3
"""Generate the golden fixtures of tests/golden/ from the REAL reference.

TEST INFRASTRUCTURE.  Runs oracle/_ref/ref_driver (the reference's own
MultipoleSubsurfaceIntegrator / BVHAccel / LayeredSkin / MPC_* code, built by
oracle/build_ref.sh from /root/reference) at --ncores 1 (deterministic order)
and packs the exported streams into compressed .npz files:

  skin_tiny.npz   scenes/skin_tiny.scene: flat BVH + meshes + lights, profile/rho tables,
                  surface points, irradiance points, octree bounds/root, camera rays,
                  closest hits, per-hit Mo, per-hit subsurface radiance, per-sample radiance
  skin_ld4.npz    scenes/skin_ld4.scene (48x27 film, 4 spp): the LDSampler stream, full Li per sample,
                  the film's image (row N1: the rest of Li, sampler and film on the device)
  tissue.npz      scenes/tissue.scene (BASELINE.json configs[0] stand-in: 160x90 film, LayeredSkin slab at the material's
                  defaults): image, octree root, strided irradiance points and per-sample radiance
  skin_lights.npz scenes/skin_lights.scene (point, sphere area, spot and distant light): light records, irradiance,
                  per-sample radiance, image
  mpc_kat.npz     MPC_ComputeDiffusionProfile on the parameter sets of the reference's
                  src/multipole/test/test.cpp:84-116 (+ thin-slab cases), raw and resampled

  skin_full.npz   scenes/skin.scene at FULL size (the benchmarked configuration): image, strided points / samples / tables
  mpc_api.npz     the MPC_* API's resample variants (default / explicit target length, arbitrary radii)
  mpc_kat512.npz  four two-layer thin-slab sets at desiredLength 512 (strided tables + both ends)

Only this script and the fixtures are committed; nothing at test time reads /root/reference.
Usage:  python tests/golden/make_golden.py [skin_tiny] [mpc_kat]
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
OUT = os.path.join(ROOT, "tests", "golden")


def read_meta(path):
    meta = {}
    for line in open(path):
        k, v = line.split()
        meta[k] = float(v) if ("." in v or "e" in v) else int(v)
    return meta


def scene_golden(name):
    scene = os.path.join(ROOT, "scenes", name + ".scene")
    with tempfile.TemporaryDirectory() as d:
        subprocess.run([DRIVER, "scene", scene, d, "--ncores", "1"], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        meta = read_meta(os.path.join(d, "meta.txt"))
        f32 = lambda fn: np.fromfile(os.path.join(d, fn), np.float32)
        i32 = lambda fn: np.fromfile(os.path.join(d, fn), np.int32)
        raw = lambda fn: np.fromfile(os.path.join(d, fn), np.uint8)
        arrays = dict(
            meta_keys=np.array(list(meta.keys())), meta_vals=np.array([float(v) for v in meta.values()]),
            bvh_nodes=raw("bvh_nodes.bin"),                      # LinearBVHNode[nNodes] 32 B
            bvh_prims=i32("bvh_prims.bin").reshape(-1, 4),      # type, mesh, index, primitiveId
            mesh_P=f32("mesh0_P.bin").reshape(-1, 3), mesh_N=f32("mesh0_N.bin").reshape(-1, 3),
            mesh_uv=f32("mesh0_uv.bin").reshape(-1, 2), mesh_idx=i32("mesh0_idx.bin").reshape(-1, 3),
            mesh_o2w=f32("mesh0_o2w.bin"),
            lights=f32("lights.bin").reshape(-1, 70),
            surface_points=raw("surface_points.bin"),            # SurfacePoint[n] 44 B
            irradiance_points=f32("irradiance_points.bin").reshape(-1, 38),
            octree_bounds=f32("octree_bounds.bin"), octree_root=f32("octree_root.bin"),
            profile_hdr=f32("profile_hdr.bin").reshape(30, 4),
            profile_rd=f32("profile_rd.bin").reshape(30, -1),
            rho_hd=f32("rho_hd.bin").reshape(-1, 30),
            skin_coeffs=f32("skin_coeffs.bin"),
            cie_xyz30=f32("cie_xyz30.bin").reshape(3, 30),
            camera_rays=f32("camera_rays.bin").reshape(-1, 8), camera_xy=f32("camera_xy.bin").reshape(-1, 2),
            hits=raw("hits.bin"),                                # DrvHit[n] 56 B
            mo=f32("mo.bin").reshape(-1, 30), L_sss=f32("L_sss.bin").reshape(-1, 30),
            L=f32("L.bin").reshape(-1, 30),
        )
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **arrays)
    print(path, os.path.getsize(path) // 1024, "KiB", {k: meta[k] for k in ("surface_points", "camera_samples", "bvh_nodes")})


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = map(int, f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4").reshape(h, w, 3)
    return data[::-1].astype(np.float32)     # PFM stores the bottom row first


def frame_golden(name):
    """Whole-frame fixture (row N1): the LDSampler stream (image position + every 1D/2D integrator
    sample of each camera sample, in task order), full Li per sample and the film's output image."""
    scene = os.path.join(ROOT, "scenes", name + ".scene")
    with tempfile.TemporaryDirectory() as d:
        subprocess.run([DRIVER, "scene", scene, d, "--ncores", "1"], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        meta = read_meta(os.path.join(d, "meta.txt"))
        f32 = lambda fn: np.fromfile(os.path.join(d, fn), np.float32)
        arrays = dict(
            meta_keys=np.array(list(meta.keys())), meta_vals=np.array([float(v) for v in meta.values()]),
            camera_xy=f32("camera_xy.bin").reshape(-1, 2),
            camera_samples=f32("camera_samples.bin").reshape(-1, meta["sample_floats"]),
            camera_rays=f32("camera_rays.bin").reshape(-1, 8),
            hits=np.fromfile(os.path.join(d, "hits.bin"), np.uint8),
            L_sss=f32("L_sss.bin").reshape(-1, 30), L=f32("L.bin").reshape(-1, 30),
            image=read_pfm(os.path.join(d, "image.pfm")),
            rgb_unit30=f32("rgb_unit30.bin"),
        )
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **arrays)
    print(path, os.path.getsize(path) // 1024, "KiB", {k: meta[k] for k in ("camera_samples", "camera_tasks", "sample_floats")})


def lights_golden(name="skin_lights"):
    """SURVEY.md 8f N4 (scenes/skin_lights.scene = .pbrt: point + sphere area + spot + distant light on the head): the
    lights as the reference built them (records in the convention of include/skin_cuda.h), every irradiance point, the full
    Li of every camera sample and the image."""
    scene = os.path.join(ROOT, "scenes", name + ".scene")
    with tempfile.TemporaryDirectory() as d:
        subprocess.run([DRIVER, "scene", scene, d, "--ncores", "1"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        meta = read_meta(os.path.join(d, "meta.txt"))
        f32 = lambda fn: np.fromfile(os.path.join(d, fn), np.float32)
        arrays = dict(
            meta_keys=np.array(list(meta.keys())), meta_vals=np.array([float(v) for v in meta.values()]),
            lights=f32("lights.bin").reshape(-1, 70),
            irradiance_E=f32("irradiance_points.bin").reshape(-1, 38)[:, 6:36].copy(),
            octree_root=f32("octree_root.bin"),
            camera_xy=f32("camera_xy.bin").reshape(-1, 2), L=f32("L.bin").reshape(-1, 30), L_sss=f32("L_sss.bin").reshape(-1, 30),
            image=read_pfm(os.path.join(d, "image.pfm")),
        )
        if os.path.exists(os.path.join(d, "mesh0_alpha.bin")):  # "texture alpha": the float image map's level-0 texels
            arrays["alpha"] = f32("mesh0_alpha.bin")
            arrays["hits"] = np.fromfile(os.path.join(d, "hits.bin"), np.uint8)
            arrays["camera_rays"] = f32("camera_rays.bin").reshape(-1, 8)
        if os.path.exists(os.path.join(d, "albedo_tex.bin")):   # "texture albedo": the MIPMap pyramid and the EWA weight table
            arrays["albedo_tex"] = f32("albedo_tex.bin")
            arrays["hits"] = np.fromfile(os.path.join(d, "hits.bin"), np.uint8)
        if os.path.exists(os.path.join(d, "bump_tex.bin")):     # "texture bumpmap": the MIPMap<float> pyramid, the bumped normals of the surface points and of the hits
            arrays["bump_tex"] = f32("bump_tex.bin")
            arrays["hits"] = np.fromfile(os.path.join(d, "hits.bin"), np.uint8)
            arrays["surface_normals"] = f32("surface_points.bin").reshape(-1, 11)[:, 3:6].copy()
        if int(meta.get("surface_points", 0)) and name == "skin_poisson":      # "bool usepoissonpointfinder": the finder's points (one task)
            arrays["surface_points"] = f32("surface_points.bin").reshape(-1, 11)
        if os.path.exists(os.path.join(d, "envmap.bin")):      # LightSource "infinite": its map and Distribution2D (ref_driver dumpLights)
            arrays["envmap"] = f32("envmap.bin")
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **arrays)
    print(path, os.path.getsize(path) // 1024, "KiB", {k: meta[k] for k in ("surface_points", "camera_samples", "lights")})


def tga_golden(name="tex_tga"):
    """scenes/skin_tga.scene: the MIPMap pyramids the reference builds from the two TGA textures (ReadImageTGA, convertIn, the
    Lanczos zoom of a 48x40 image to 64x64, the box-filtered levels) -- what the host mirror has to hand to the device."""
    arrays = {}
    for suffix, scene_name in (("", "skin_tga.scene"), ("_b", "skin_tga_b.scene")):      # _b: the other TGA encodings (tools/make_tga_textures.py)
        scene = os.path.join(ROOT, "scenes", scene_name)
        with tempfile.TemporaryDirectory() as d:
            subprocess.run([DRIVER, "scene", scene, d, "--ncores", "1"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            arrays["albedo_tex" + suffix] = np.fromfile(os.path.join(d, "albedo_tex.bin"), np.float32)
            arrays["bump_tex" + suffix] = np.fromfile(os.path.join(d, "bump_tex.bin"), np.float32)
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **arrays)
    print(path, os.path.getsize(path) // 1024, "KiB")


def tissue_golden(name="tissue"):
    """BASELINE.json configs[0] stand-in (scenes/tissue.scene = scenes/tissue.pbrt: LayeredSkin slab at the material's
    defaults incl. the transmission lobe, mesh without normals / uvs): image, octree root and bounds, and every 37th
    irradiance point / every 29th camera sample's radiance, to keep the fixture small."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import make_slab_mesh
    g = os.path.join(ROOT, "scenes", "geometry")
    make_slab_mesh.write(os.path.join(g, "slab.mesh.bin"), os.path.join(g, "slab.pbrt"))
    scene = os.path.join(ROOT, "scenes", name + ".scene")
    with tempfile.TemporaryDirectory() as d:
        subprocess.run([DRIVER, "scene", scene, d, "--ncores", "1"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        meta = read_meta(os.path.join(d, "meta.txt"))
        f32 = lambda fn: np.fromfile(os.path.join(d, fn), np.float32)
        irr = f32("irradiance_points.bin").reshape(-1, 38)
        L = f32("L.bin").reshape(-1, 30)
        arrays = dict(
            meta_keys=np.array(list(meta.keys())), meta_vals=np.array([float(v) for v in meta.values()]),
            irradiance_stride=np.int64(37), irradiance_points=irr[::37].copy(),
            octree_bounds=f32("octree_bounds.bin"), octree_root=f32("octree_root.bin"),
            L_stride=np.int64(29), L=L[::29].copy(), L_sss=f32("L_sss.bin").reshape(-1, 30)[::29].copy(),
            camera_xy=f32("camera_xy.bin").reshape(-1, 2)[::29].copy(),
            image=read_pfm(os.path.join(d, "image.pfm")),
        )
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **arrays)
    print(path, os.path.getsize(path) // 1024, "KiB", {k: meta[k] for k in ("surface_points", "camera_samples", "bvh_nodes")})


# (nLayers, desiredLength, lerpOnThinSlab, resample, stepSize, [(ior, thickness, mua, musp)...])
# First three: src/multipole/test/test.cpp:84-116 at desiredLength 512 -> here 64/128 to keep
# the fixture small; the full-size known answers are checked in the GPU-marked tests against
# oracle/_ref directly when it is present.
EPI = (1.4, 0.025, 0.268088, 19.4879)
DERM = (1.4, 2.0, 0.268088, 9.74395)
MPC_CASES = [
    (1, 64, 1, 0, 0.0, [EPI]),
    (1, 64, 1, 0, 0.0, [DERM]),
    (2, 64, 1, 0, 0.0, [EPI, DERM]),
    (2, 64, 1, 1, 0.0, [EPI, DERM]),
    (2, 128, 1, 1, 0.0, [EPI, DERM]),
    (2, 64, 0, 1, 0.0, [EPI, DERM]),
    (2, 64, 1, 1, 0.0, [(1.4, 0.005, 0.5, 30.0), (1.3, 1.0, 0.1, 5.0)]),
    (3, 64, 1, 1, 0.0, [(1.4, 0.01, 0.5, 30.0), (1.45, 0.1, 0.3, 15.0), (1.35, 3.0, 0.05, 8.0)]),
]


def mpc_golden():
    with tempfile.TemporaryDirectory() as d:
        inp, outp = os.path.join(d, "in.txt"), os.path.join(d, "out.bin")
        with open(inp, "w") as f:
            for nl, dl, lerp, res, step, layers in MPC_CASES:
                f.write("%d %d %d %d %r " % (nl, dl, lerp, res, step) + " ".join("%r %r %r %r" % l for l in layers) + "\n")
        subprocess.run([DRIVER, "mpc", inp, outp], check=True, stdout=subprocess.DEVNULL)
        blob = open(outp, "rb").read()
    arrays, off = {}, 0
    for i, case in enumerate(MPC_CASES):
        ln = int(np.frombuffer(blob, np.uint32, 1, off)[0]); off += 4
        tot = np.frombuffer(blob, np.float32, 2, off); off += 8
        body = np.frombuffer(blob, np.float32, 3 * ln, off).reshape(3, ln); off += 12 * ln
        arrays["case%d_spec" % i] = np.array([case[0], case[1], case[2], case[3], case[4]] + [x for l in case[5] for x in l], np.float64)
        arrays["case%d_totals" % i] = tot.copy()
        arrays["case%d_d2_R_T" % i] = body.copy()
    path = os.path.join(OUT, "mpc_kat.npz")
    np.savez_compressed(path, **arrays)
    print(path, os.path.getsize(path) // 1024, "KiB")


def strided_table(a, stride=16, edge=2048):
    """A long table row(s) [..., len] kept as: every stride-th entry, the first `edge` and the last `edge` entries."""
    a = np.asarray(a)
    return dict(strided=a[..., ::stride].copy(), head=a[..., :edge].copy(), tail=a[..., -edge:].copy())


SKIN_FULL_FLIPS = [3311269, 19334839]


def skin_full_golden(name="skin_full"):
    """BASELINE.json configs[1] at FULL size (scenes/skin.scene: 24.6 M irradiance points, 1280x720x4 spp, desiredLength 512):
    the reference's image, octree root and bounds, every 4099th irradiance point, every 997th camera sample (hit, Mo,
    subsurface radiance, full Li) and the 30 profile tables (every 16th entry + both ends) with their headers.
    The run takes ~22 min at --ncores 1 and ~9 GB of scratch; SKIN_FULL_DIR=<dir> reuses an existing export."""
    d = os.environ.get("SKIN_FULL_DIR")
    tmp = None
    if not d:
        tmp = tempfile.TemporaryDirectory(); d = tmp.name
        subprocess.run([DRIVER, "scene", os.path.join(ROOT, "scenes", "skin.scene"), d, "--ncores", "1"], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    meta = read_meta(os.path.join(d, "meta.txt"))
    mm = lambda fn, dt, cols: np.memmap(os.path.join(d, fn), dt, "r").reshape(-1, cols)
    irr = mm("irradiance_points.bin", np.float32, 38)
    si, sc = 4099, 997
    hits = np.memmap(os.path.join(d, "hits.bin"), np.uint8, "r").reshape(-1, 56)
    rd = mm("profile_rd.bin", np.float32, meta["profile_len"])
    tab = strided_table(rd)
    arrays = dict(
        meta_keys=np.array(list(meta.keys())), meta_vals=np.array([float(v) for v in meta.values()]),
        irradiance_stride=np.int64(si), irradiance_points=np.array(irr[::si]),
        octree_bounds=np.fromfile(os.path.join(d, "octree_bounds.bin"), np.float32), octree_root=np.fromfile(os.path.join(d, "octree_root.bin"), np.float32),
        sample_stride=np.int64(sc), camera_xy=np.array(mm("camera_xy.bin", np.float32, 2)[::sc]), camera_rays=np.array(mm("camera_rays.bin", np.float32, 8)[::sc]),
        hits=np.array(hits[::sc]), mo=np.array(mm("mo.bin", np.float32, 30)[::sc]), L_sss=np.array(mm("L_sss.bin", np.float32, 30)[::sc]),
        L=np.array(mm("L.bin", np.float32, 30)[::sc]),
        profile_hdr=np.fromfile(os.path.join(d, "profile_hdr.bin"), np.float32).reshape(30, 4),
        profile_stride=np.int64(16), profile_rd_strided=tab["strided"], profile_rd_head=tab["head"], profile_rd_tail=tab["tail"],
        rho_hd=np.fromfile(os.path.join(d, "rho_hd.bin"), np.float32).reshape(-1, 30)[:1025],
        image=read_pfm(os.path.join(d, "image.pfm")),
        # the reference's rows of the irradiance points where the device pass is known to differ by more than 1e-4 (found by
        # tools/diag_irradiance_full.py over all 24 648 458 points: ONE shadow-ray visibility flip, one point off by 8e-4)
        irradiance_flip_idx=np.array(SKIN_FULL_FLIPS, np.int64), irradiance_flip_rows=np.array(irr[SKIN_FULL_FLIPS]),
    )
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **arrays)
    print(path, os.path.getsize(path) // 1024, "KiB", {k: meta[k] for k in ("surface_points", "camera_samples", "profile_len")})
    if tmp:
        tmp.cleanup()


# BASELINE.json configs[4] at the size that is benchmarked: two-layer thin-epidermis sets at desiredLength 512
# (2048^2 transforms, 119 766-entry tables).  Epidermis thickness as a multiple of its mean free path: 0.02, 0.5, 1.2, 2.0.
C5_EPI_MFP = 1.0 / (0.268088 + 19.4879)
MPC512_CASES = [(2, 512, 1, 1, 0.0, [(1.4, t * C5_EPI_MFP, 0.268088, 19.4879), (1.4, 2.0, 0.268088, 9.74395)]) for t in (0.02, 0.5, 1.2, 2.0)]


# The MPC_* DLL API as its three call sites use it (core/multipole.cpp:268-291, renderers/mcprofile.cpp:402-423,
# renderers/profilefit.cpp:190-196): default-target and explicit-target uniform resampling, and MPC_ResampleDistribution at
# arbitrary radii (ref_driver mpc resample codes 1, 3, 2).
MPC_API_CASES = [
    (2, 64, 1, 1, 0.0, [EPI, DERM]),
    (2, 64, 1, 3, 0.0, [EPI, DERM]),
    (2, 64, 1, 2, 0.0, [EPI, DERM]),
    (1, 128, 0, 2, 0.0, [DERM]),
    (3, 64, 1, 0, 0.0, [(1.4, 0.01, 0.5, 30.0), (1.45, 0.1, 0.3, 15.0), (1.35, 3.0, 0.05, 8.0)]),
]


def mpc_api_golden():
    with tempfile.TemporaryDirectory() as d:
        inp, outp = os.path.join(d, "in.txt"), os.path.join(d, "out.bin")
        with open(inp, "w") as f:
            for nl, dl, lerp, res, step, layers in MPC_API_CASES:
                f.write("%d %d %d %d %r " % (nl, dl, lerp, res, step) + " ".join("%r %r %r %r" % l for l in layers) + "\n")
        subprocess.run([DRIVER, "mpc", inp, outp], check=True, stdout=subprocess.DEVNULL)
        blob = open(outp, "rb").read()
    arrays, off = {}, 0
    for i, case in enumerate(MPC_API_CASES):
        ln = int(np.frombuffer(blob, np.uint32, 1, off)[0]); off += 4
        tot = np.frombuffer(blob, np.float32, 2, off); off += 8
        body = np.frombuffer(blob, np.float32, 3 * ln, off).reshape(3, ln); off += 12 * ln
        arrays["case%d_spec" % i] = np.array([case[0], case[1], case[2], case[3], case[4]] + [x for l in case[5] for x in l], np.float64)
        arrays["case%d_totals" % i] = tot.copy()
        arrays["case%d_d2_R_T" % i] = body.copy()
    path = os.path.join(OUT, "mpc_api.npz")
    np.savez_compressed(path, **arrays)
    print(path, os.path.getsize(path) // 1024, "KiB")


def mpc512_golden():
    with tempfile.TemporaryDirectory() as d:
        inp, outp = os.path.join(d, "in.txt"), os.path.join(d, "out.bin")
        with open(inp, "w") as f:
            for nl, dl, lerp, res, step, layers in MPC512_CASES:
                f.write("%d %d %d %d %r " % (nl, dl, lerp, res, step) + " ".join("%r %r %r %r" % l for l in layers) + "\n")
        subprocess.run([DRIVER, "mpc", inp, outp], check=True, stdout=subprocess.DEVNULL)
        blob = open(outp, "rb").read()
    arrays, off = {}, 0
    for i, case in enumerate(MPC512_CASES):
        ln = int(np.frombuffer(blob, np.uint32, 1, off)[0]); off += 4
        tot = np.frombuffer(blob, np.float32, 2, off); off += 8
        body = np.frombuffer(blob, np.float32, 3 * ln, off).reshape(3, ln); off += 12 * ln
        arrays["case%d_spec" % i] = np.array([case[0], case[1], case[2], case[3], case[4]] + [x for l in case[5] for x in l], np.float64)
        arrays["case%d_totals" % i] = tot.copy(); arrays["case%d_len" % i] = np.int64(ln)
        for k, v in strided_table(body).items():
            arrays["case%d_d2_R_T_%s" % (i, k)] = v
    path = os.path.join(OUT, "mpc_kat512.npz")
    np.savez_compressed(path, **arrays)
    print(path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    what = sys.argv[1:] or ["skin_tiny", "skin_ld4", "tissue", "skin_lights", "skin_albedo", "mpc_kat"]
    if not os.path.exists(DRIVER):
        sys.exit("oracle/_ref/ref_driver missing: run oracle/build_ref.sh (needs /root/reference)")
    for w in what:
        mpc_golden() if w == "mpc_kat" else mpc512_golden() if w == "mpc_kat512" else mpc_api_golden() if w == "mpc_api" else tga_golden(w) if w == "tex_tga" else skin_full_golden(w) if w == "skin_full" else frame_golden(w) if w == "skin_ld4" else tissue_golden(w) if w == "tissue" else lights_golden(w) if w in ("skin_lights", "skin_albedo", "skin_env", "skin_alpha", "skin_tex", "skin_bump", "skin_poisson") else scene_golden(w)
