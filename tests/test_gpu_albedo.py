"""LayeredSkin's constant "color albedo" with a non-default `mix` (materials/layeredskin.cpp:183,247): the irradiance pass scales
E by Pow(albedo, mix) (integrators/multipolesubsurface.cpp:139-140) and Li scales the subsurface term by Pow(albedo, 1 - mix)
(:287).  Device path against the real reference's export of scenes/skin_albedo.scene (tests/golden/skin_albedo.npz,
oracle/_ref/ref_driver --ncores 1).  The two 30-entry factors are made on the host with libm's powf (host/skin_render.cpp
InstallAlbedo) and installed with skc_set_albedo_factors.
Bars (BASELINE.json): irradiance 1e-4 relative per point; image RMSE 1e-3 relative per channel."""
import os

import numpy as np
import pytest

import golden_util as gu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def golden():
    z = np.load(os.path.join(ROOT, "tests", "golden", "skin_albedo.npz"))
    return z, dict(zip([str(k) for k in z["meta_keys"]], z["meta_vals"]))


def test_pbrt_file_equals_line_format(pkg):
    """`"color albedo"` and `"float mix"` through the .pbrt front end give the material and integrator of the .scene file (CPU)."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import mesh_bin_to_pbrt
    dst = os.path.join(ROOT, "scenes", "geometry", "head.pbrt")
    if not os.path.exists(dst):
        mesh_bin_to_pbrt.convert(os.path.join(ROOT, "scenes", "head.mesh.bin"), dst)
    a = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_albedo.scene")); b = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_albedo.pbrt"))
    try:
        for k in ("nodes", "prims", "lights", "P", "idx", "o2w", "skin"):
            assert np.array_equal(np.ascontiguousarray(getattr(a, k)).view(np.uint8), np.ascontiguousarray(getattr(b, k)).view(np.uint8)), k
        assert a.mix == b.mix == np.float32(0.3)
        assert np.array_equal(a.albedo_rgb(), b.albedo_rgb()) and list(a.albedo_rgb()) == [np.float32(0.8), np.float32(0.55), np.float32(0.45)]
    finally:
        a.close(); b.close()


@pytest.mark.gpu
def test_irradiance_and_frame_with_albedo(pkg, golden):
    z, meta = golden
    dev = pkg.SkinDevice(0)         # its own context: the albedo factors stay installed on a context until they are replaced
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_albedo.scene"))
    try:
        assert abs(float(meta["mix"]) - 0.3) < 1e-7
        hs.upload(dev)
        hs.build_profile(dev)       # installs the factors (skh_build_profile -> InstallAlbedo)
        n_pts = dev.tessellate_resident(hs.minsampledistance, False, material_id=int(meta["skin_material_id"]))
        assert n_pts == int(meta["surface_points"])
        dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
        _, E = dev.irradiance_rows(0, n_pts, 1)
        ref = z["irradiance_E"]
        e = gu.row_rel_err(E, ref)
        lit = ref.max(axis=1) > 0
        print("irradiance with albedo^mix: %d points (%d lit), bit-exact rows %.4f, max rel %.3g" % (len(ref), lit.sum(), np.mean(e == 0), e.max()))
        assert np.array_equal(E.max(axis=1) > 0, lit)
        assert e.max() < 1e-6, e.max()
        # it is a parameter of the context, not baked in: cleared, the rows are the albedo-1 rows -- larger by 1 / Pow(albedo, mix),
        # 1 / 0.8^0.3 = 1.07 at the red end, 1 / 0.45^0.3 = 1.27 at the blue end -- and installing it again brings the reference's rows back
        dev.set_albedo_factors(None, None)
        dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
        _, E1 = dev.irradiance_rows(0, n_pts, 1)
        k = int(np.argmax(ref[:, 29]))
        assert 1.05 * ref[k, 29] < E1[k, 29] < 1.10 * ref[k, 29] and 1.2 * ref[k, 0] < E1[k, 0] < 1.35 * ref[k, 0]
        hs.build_profile(dev)
        dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
        _, E2 = dev.irradiance_rows(0, n_pts, 1)
        assert np.array_equal(E2.view(np.uint32), E.view(np.uint32))
        _, root, _ = dev.build_octree()
        er = np.abs(root - z["octree_root"]) / np.maximum(np.abs(z["octree_root"]), 1e-30)
        assert er.max() < 1e-5
        n = dev.frame_setup(hs.frame_desc(n_cores=1))
        assert n == int(meta["camera_samples"])
        dev.frame_samples()
        dev.frame_film_clear()
        Lr = dev.frame_render(0, n, hs.maxerror, hs.mix, want_L=True)
        order = dev.frame_order(0, n)
        L = np.empty_like(Lr); L[order] = Lr
        ref = z["L"]
        lit = ref.max(axis=1) > 0
        assert np.array_equal(L.max(axis=1) > 0, lit)
        e = gu.row_rel_err(L, ref)[lit]
        print("full Li with albedo^(1-mix): %d lit samples, max rel %.3g" % (lit.sum(), e.max()))
        assert e.max() < 1e-5, e.max()
        rgb = hs.film_to_rgb(dev.frame_film())
        for c in range(3):
            rmse = np.sqrt(np.mean((rgb[..., c] - z["image"][..., c]) ** 2)); s = np.sqrt(np.mean(z["image"][..., c] ** 2))
            print("channel %d: RMSE / RMS = %.3g" % (c, rmse / s))
            assert rmse < 1e-3 * s
    finally:
        hs.close()
        dev.close()
