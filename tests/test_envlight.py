"""SURVEY.md 8f N4, LightSource "infinite" (lights/infinite.cpp): the head under a 16 x 8 environment map (scenes/env16x8.pfm,
tools/make_envmap.py) rotated about z, plus a sphere emitter -- scenes/skin_env.scene, against the real reference's export
(tests/golden/skin_env.npz, oracle/_ref/ref_driver --ncores 1).

CPU: the host mirror's InfiniteAreaLight construction (host/skin_envmap.cpp: PFM reader, MIPMap pyramid and trilinear lookup,
Distribution2D) gives the reference's texels, pdfs and cdfs bit for bit.  GPU: irradiance (InfiniteAreaLight::Sample_L), the
complete Li (both MIS terms of EstimateDirect with Pdf and Le for rays that leave the scene, Le for camera rays that miss) and
the image.  Bars (BASELINE.json): irradiance 1e-4 per point, image RMSE 1e-3 per channel."""
import os

import numpy as np
import pytest

import golden_util as gu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def golden():
    z = np.load(os.path.join(ROOT, "tests", "golden", "skin_env.npz"))
    return z, dict(zip([str(k) for k in z["meta_keys"]], z["meta_vals"]))


def _split_env(flat):
    w, h = [int(x) for x in flat[:2].view(np.int32)]
    sizes = [("rgb", 3 * w * h), ("condFunc", w * h), ("condCdf", h * (w + 1)), ("condInt", h), ("margFunc", h), ("margCdf", h + 1), ("margInt", 1)]
    out, o = {"w": w, "h": h}, 2
    for k, n in sizes:
        out[k] = flat[o:o + n]; o += n
    assert o == len(flat)
    return out


@pytest.mark.parametrize("ext", ["scene", "pbrt"])
def test_infinite_light_construction_bit_exact(pkg, golden, ext):
    z, _ = golden
    if ext == "pbrt":
        import sys
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import mesh_bin_to_pbrt
        dst = os.path.join(ROOT, "scenes", "geometry", "head.pbrt")
        if not os.path.exists(dst):
            mesh_bin_to_pbrt.convert(os.path.join(ROOT, "scenes", "head.mesh.bin"), dst)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_env." + ext))
    try:
        assert hs.lights.shape == z["lights"].shape == (2, 70)
        assert np.array_equal(hs.lights.view(np.uint32), z["lights"].view(np.uint32)), "light records (LightToWorld of the rotated map, map size, nSamples)"
        assert list(hs.lights[:, 32]) == [-4, np.float32(0.2)] and list(hs.light_samples()) == [8, 4]
        ref, got = _split_env(z["envmap"]), hs.envmap()
        assert (got["w"], got["h"]) == (ref["w"], ref["h"]) == (16, 8)
        for k in ("rgb", "condFunc", "condCdf", "condInt", "margFunc", "margCdf"):
            assert np.array_equal(got[k].view(np.uint32), ref[k].view(np.uint32)), k
        assert np.float32(got["margInt"]).view(np.uint32) == ref["margInt"].view(np.uint32)[0]
    finally:
        hs.close()


def test_constant_infinite_light_is_a_one_texel_map(pkg, tmp_path):
    src = open(os.path.join(ROOT, "scenes", "skin_env.scene")).read().replace("envlight env16x8.pfm", "envlight none")
    src = src.replace("mesh head.mesh.bin", "mesh %s" % os.path.join(ROOT, "scenes", "head.mesh.bin"))
    (tmp_path / "c.scene").write_text(src)
    hs = pkg.HostScene(str(tmp_path / "c.scene"))
    try:
        e = hs.envmap()
        assert (e["w"], e["h"]) == (1, 1) and list(e["condCdf"]) == [0, 1] and list(e["margCdf"]) == [0, 1]
        assert np.all(e["rgb"] > 1.0) and e["rgb"][0] > e["rgb"][2], "L = (1.5, 1.4, 1.2) through FromRGB / ToRGB"
    finally:
        hs.close()


def test_non_power_of_two_map_is_refused(pkg, tmp_path):
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import make_envmap
    make_envmap.write_pfm(str(tmp_path / "odd.pfm"), make_envmap.envmap(12, 8))
    src = open(os.path.join(ROOT, "scenes", "skin_env.scene")).read().replace("envlight env16x8.pfm", "envlight %s" % (tmp_path / "odd.pfm"))
    src = src.replace("mesh head.mesh.bin", "mesh %s" % os.path.join(ROOT, "scenes", "head.mesh.bin"))
    (tmp_path / "o.scene").write_text(src)
    with pytest.raises(Exception, match="power of two"):
        pkg.HostScene(str(tmp_path / "o.scene"))


@pytest.mark.gpu
def test_device_acosf_atan2f_equal_libm(dev):
    """InfiniteAreaLight::Pdf / Le go through SphericalTheta / SphericalPhi = acosf / atan2f (core/geometry.h); the device restates
    glibc's (csrc/skc_common.cuh; tools/libm_check.c checked them over the whole domain / 4e8 pairs on the CPU)."""
    import ctypes
    libm = ctypes.CDLL("libm.so.6")
    rng = np.random.default_rng(11)
    x = np.concatenate([rng.uniform(-1, 1, 150000), 1 - rng.random(30000) ** 6, -1 + rng.random(30000) ** 6, rng.random(20000) ** 8 - 0.5 * 0, [0.0, 0.5, -0.5, 1.0, -1.0, 2.0 ** -30]]).astype(np.float32)
    f = libm.acosf; f.restype = ctypes.c_float; f.argtypes = [ctypes.c_float]
    host = np.array([f(float(v)) for v in x], np.float32)
    assert np.array_equal(dev.test_libmf(5, x).view(np.uint32), host.view(np.uint32))
    ang = rng.uniform(0, 2 * np.pi, 200000); r = np.exp(rng.uniform(-8, 8, 200000))
    y = (r * np.sin(ang)).astype(np.float32); xx = (r * np.cos(ang)).astype(np.float32)
    y[:200] = 0; xx[200:400] = 0; xx[400:600] = 1
    pairs = np.concatenate([y, xx[::-1]]).astype(np.float32)          # kernel: atan2f(v[i], v[n - 1 - i])
    g = libm.atan2f; g.restype = ctypes.c_float; g.argtypes = [ctypes.c_float, ctypes.c_float]
    n = len(pairs)
    host = np.array([g(float(pairs[i]), float(pairs[n - 1 - i])) for i in range(n)], np.float32)
    assert np.array_equal(dev.test_libmf(6, pairs).view(np.uint32), host.view(np.uint32))


@pytest.mark.gpu
def test_irradiance_and_frame_under_the_infinite_light(pkg, golden):
    z, meta = golden
    dev = pkg.SkinDevice(0)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_env.scene"))
    try:
        hs.upload(dev)
        hs.build_profile(dev)
        n_pts = dev.tessellate_resident(hs.minsampledistance, False, material_id=int(meta["skin_material_id"]))
        assert n_pts == int(meta["surface_points"])
        dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
        _, E = dev.irradiance_rows(0, n_pts, 1)
        ref = z["irradiance_E"]
        e = gu.row_rel_err(E, ref)
        lit = ref.max(axis=1) > 0
        print("irradiance under the environment map + 1 sphere: %d points (%d lit), bit-exact rows %.4f, max rel %.3g" % (len(ref), lit.sum(), np.mean(e == 0), e.max()))
        assert np.array_equal(E.max(axis=1) > 0, lit)
        assert e.max() < 1e-6, e.max()
        _, root, _ = dev.build_octree()
        er = np.abs(root - z["octree_root"]) / np.maximum(np.abs(z["octree_root"]), 1e-30)
        assert er.max() < 1e-5
        # ---- the frame: every camera sample carries radiance now (the ones that miss see the map)
        n = dev.frame_setup(hs.frame_desc(n_cores=1))
        assert n == int(meta["camera_samples"])
        dev.frame_samples()
        xy, _, _ = dev.frame_get_samples(0, n)
        assert np.array_equal(xy.view(np.uint32), z["camera_xy"].view(np.uint32)), "the sampler stream with an 8-sample and a 4-sample light"
        dev.frame_film_clear()
        Lr = dev.frame_render(0, n, hs.maxerror, hs.mix, want_L=True)
        order = dev.frame_order(0, n)
        L = np.empty_like(Lr); L[order] = Lr
        ref = z["L"]
        lit = ref.max(axis=1) > 0
        assert lit.mean() > 0.95 and np.array_equal(L.max(axis=1) > 0, lit)
        e = gu.row_rel_err(L, ref)
        miss = z["L_sss"].max(axis=1) == 0
        print("full Li under the environment map: %d samples, max rel %.3g (rows without a subsurface term -- background and emitter: %d, bit-exact %.4f)"
              % (len(ref), e.max(), miss.sum(), np.mean(e[miss] == 0)))
        assert e.max() < 1e-5, e.max()
        rgb = hs.film_to_rgb(dev.frame_film())
        for c in range(3):
            rmse = np.sqrt(np.mean((rgb[..., c] - z["image"][..., c]) ** 2)); s = np.sqrt(np.mean(z["image"][..., c] ** 2))
            print("channel %d: RMSE / RMS = %.3g" % (c, rmse / s))
            assert rmse < 1e-3 * s
    finally:
        hs.close()
        dev.close()
