"""LightSource "point", "spot" and "distant" next to a sphere area light, in the irradiance pass
(IrradianceTask::Run, integrators/multipolesubsurface.cpp:112-136 -> Light::Sample_L of lights/point.cpp:49-56, spot.cpp:52-72,
distant.cpp:49-56) and in direct lighting (EstimateDirect's delta-light branch, core/integrator.cpp:114-134), against the real
reference's export of scenes/skin_lights.scene (tests/golden/skin_lights.npz, oracle/_ref/ref_driver --ncores 1).
Bars (BASELINE.json): irradiance 1e-4 relative per point; image RMSE 1e-3 relative per channel; light records bit-exact."""
import os

import numpy as np
import pytest

import golden_util as gu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def golden():
    z = np.load(os.path.join(ROOT, "tests", "golden", "skin_lights.npz"))
    return z, dict(zip([str(k) for k in z["meta_keys"]], z["meta_vals"]))


@pytest.mark.parametrize("ext", ["scene", "pbrt"])
def test_light_records_bit_exact(pkg, golden, ext):
    """CreatePointLight / CreateSpotLight / CreateDistantLight in the host mirror: LightToWorld and its inverse, cone cosines,
    direction and the FromRGB spectra equal the reference's objects bit for bit (CPU test)."""
    z, _ = golden
    if ext == "pbrt":
        import sys
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import mesh_bin_to_pbrt
        dst = os.path.join(ROOT, "scenes", "geometry", "head.pbrt")
        if not os.path.exists(dst):
            mesh_bin_to_pbrt.convert(os.path.join(ROOT, "scenes", "head.mesh.bin"), dst)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_lights." + ext))
    assert hs.lights.shape == z["lights"].shape == (4, 70)
    assert np.array_equal(hs.lights.view(np.uint32), z["lights"].view(np.uint32))
    assert list(hs.lights[:, 32]) == [-1, np.float32(0.2), -2, -3] and list(hs.light_samples()) == [1, 4, 1, 1]
    assert hs.n_prims == hs.n_tris + 1, "only the sphere emitter is a primitive; delta lights have no shape"


def test_coloured_rgb_to_spectrum(pkg, tmp_path):
    """Spectrum::FromRGB for non-grey values (core/spectrum.cpp:103-145): the three branches pick the right basis pair, the
    result is non-negative, and a grey value reduces to the white basis."""
    src = open(os.path.join(ROOT, "scenes", "skin_lights.scene")).read().replace("pointlight 1.5 -3 1.5 30 40 55", "pointlight 1.5 -3 1.5 40 10 25")
    src = src.replace("mesh head.mesh.bin", "mesh %s" % os.path.join(ROOT, "scenes", "head.mesh.bin"))
    (tmp_path / "c.scene").write_text(src)
    hs = pkg.HostScene(str(tmp_path / "c.scene"))
    I = hs.lights[0, 40:]
    assert np.all(I >= 0) and I[-1] > I[15] and I[0] > I[15], "red and blue ends above the green middle for rgb (40, 10, 25)"
    assert 9 < I[15] < 16 and 30 < I[-1] < 42


@pytest.mark.gpu
def test_irradiance_and_frame_under_delta_lights(pkg, dev, golden):
    z, meta = golden
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_lights.scene"))
    try:
        hs.upload(dev)
        hs.build_profile(dev)
        n_pts = dev.tessellate_resident(hs.minsampledistance, False, material_id=int(meta["skin_material_id"]))
        assert n_pts == int(meta["surface_points"])
        n_tasks = hs.irradiance_tasks(n_pts, 1)
        irr = dev.irradiance(dev.get_surface_points(0, n_pts), n_tasks, hs.mix)
        ref = z["irradiance_E"]
        e = gu.row_rel_err(irr["E"], ref)
        lit = ref.max(axis=1) > 0
        print("irradiance under 4 lights: %d points (%d lit), within 1e-6: %.4f, max rel %.3g" % (len(ref), lit.sum(), np.mean(e < 1e-6), e.max()))
        assert np.array_equal(irr["E"].max(axis=1) > 0, lit)
        assert e.max() < 1e-6, e.max()
        _, root, _ = dev.build_octree()
        er = np.abs(root - z["octree_root"]) / np.maximum(np.abs(z["octree_root"]), 1e-30)
        assert er.max() < 1e-4
        n = dev.frame_setup(hs.frame_desc(n_cores=1))
        assert n == int(meta["camera_samples"])
        dev.frame_samples()
        xy, _, _ = dev.frame_get_samples(0, n)
        assert np.array_equal(xy.view(np.uint32), z["camera_xy"].view(np.uint32)), "the sampler stream with 1-sample light arrays"
        dev.frame_film_clear()
        Lr = dev.frame_render(0, n, hs.maxerror, hs.mix, want_L=True)
        order = dev.frame_order(0, n)
        L = np.empty_like(Lr); L[order] = Lr
        ref = z["L"]
        lit = ref.max(axis=1) > 0
        assert np.array_equal(L.max(axis=1) > 0, lit)
        e = gu.row_rel_err(L, ref)[lit]
        print("full Li: %d lit samples, within 1e-4: %.4f, within 1e-2: %.4f, max %.3g" % (lit.sum(), np.mean(e < 1e-4), np.mean(e < 1e-2), e.max()))
        assert e.max() < 1e-5, e.max()
        d_ref = ref - z["L_sss"]; d_got = L - z["L_sss"]
        scale = np.abs(d_ref).max(axis=1)
        m = scale > 1e-3 * np.abs(ref).max(axis=1)
        ed = np.abs(d_got - d_ref).max(axis=1)[m] / scale[m]
        print("direct term alone: %d samples, median rel %.3g, within 1e-3: %.4f" % (m.sum(), np.median(ed), np.mean(ed < 1e-3)))
        assert ed.max() < 1e-3, ed.max()     # a difference of two nearly equal rows amplifies the rounding of the sum
        rgb = hs.film_to_rgb(dev.frame_film())
        for c in range(3):
            rmse = np.sqrt(np.mean((rgb[..., c] - z["image"][..., c]) ** 2)); s = np.sqrt(np.mean(z["image"][..., c] ** 2))
            print("channel %d: RMSE / RMS = %.3g" % (c, rmse / s))
            assert rmse < 1e-3 * s
    finally:
        hs.close()
