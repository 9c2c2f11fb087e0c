"""BASELINE.json configs[0]: the tissue slab at reduced resolution (scenes/tissue.scene = scenes/tissue.pbrt, the stand-in
for the scenes/tissue.pbrt the fork does not ship -- SURVEY.md 8d C1), device path against the real reference's export
(tests/golden/tissue.npz from oracle/_ref/ref_driver at --ncores 1).  What this scene adds over the head scenes: the
LayeredSkin material at its own defaults, i.e. with the MicrofacetTransmission lobe on (Kt = 1), a mesh without vertex
normals or uvs (TriangleBase::GetShadingGeometry returns the geometric frame, shapes/trianglemesh.inl:227-230), an
emitter close to the surface, and an open two-sided slab.

Tolerances (BASELINE.json): irradiance 1e-4 relative per point, image RMSE 1e-3 relative per channel; per-sample radiance
1e-4 for the bulk of the samples (a shadow ray that differs in the last ulp can flip one of a sample's lighting terms)."""
import os
import sys

import numpy as np
import pytest

import golden_util as gu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def slab():
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import make_slab_mesh
    g = os.path.join(ROOT, "scenes", "geometry")
    if not (os.path.exists(os.path.join(g, "slab.mesh.bin")) and os.path.exists(os.path.join(g, "slab.pbrt"))):
        make_slab_mesh.write(os.path.join(g, "slab.mesh.bin"), os.path.join(g, "slab.pbrt"))


@pytest.fixture(scope="module")
def golden():
    z = np.load(os.path.join(ROOT, "tests", "golden", "tissue.npz"))
    return z, dict(zip([str(k) for k in z["meta_keys"]], z["meta_vals"]))


def test_pbrt_file_equals_line_format(pkg, slab):
    a = pkg.HostScene(os.path.join(ROOT, "scenes", "tissue.scene")); b = pkg.HostScene(os.path.join(ROOT, "scenes", "tissue.pbrt"))
    for k in ("nodes", "prims", "lights", "P", "idx", "o2w", "skin"):
        assert np.array_equal(np.ascontiguousarray(getattr(a, k)).view(np.uint8), np.ascontiguousarray(getattr(b, k)).view(np.uint8)), k
    assert a.N is None and b.N is None and a.uv is None and b.uv is None
    assert np.array_equal(a.frame_desc(1), b.frame_desc(1))


def test_tissue_preprocess_and_frame(pkg, dev, slab, golden):
    z, meta = golden
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "tissue.scene"))
    try:
        assert hs.n_nodes == int(meta["bvh_nodes"]) and hs.n_prims == int(meta["bvh_prims"])
        hs.upload(dev)
        hs.build_profile(dev)
        n_pts = dev.tessellate_resident(hs.minsampledistance, False, material_id=int(meta["skin_material_id"]))
        assert n_pts == int(meta["surface_points"])
        n_tasks = hs.irradiance_tasks(n_pts, 1)
        assert n_tasks == int(meta["irradiance_tasks"])
        irr = dev.irradiance(dev.get_surface_points(0, n_pts), n_tasks, hs.mix)     # device-tessellated points through the host-buffer call: records come back
        ref = np.ascontiguousarray(z["irradiance_points"]).view(gu.DT_IP).reshape(-1)
        got = irr[::int(z["irradiance_stride"])]
        assert np.array_equal(got["p"].view(np.uint32), ref["p"].view(np.uint32)) and np.array_equal(got["area"].view(np.uint32), ref["area"].view(np.uint32))
        e = gu.row_rel_err(got["E"], ref["E"])
        lit = ref["E"].max(axis=1) > 0
        print("irradiance: %d sampled points (%d lit), max rel %.3g, bit-exact rows %.4f" % (len(ref), lit.sum(), e.max(), np.mean((got["E"] == ref["E"]).all(axis=1))))
        assert np.array_equal(got["E"].max(axis=1) > 0, lit) and e.max() < 1e-4
        _, root, bounds = dev.build_octree()
        assert np.array_equal(bounds.view(np.uint32), z["octree_bounds"].view(np.uint32))
        er = np.abs(root - z["octree_root"]) / np.maximum(np.abs(z["octree_root"]), 1e-30)
        print("octree root max rel %.3g" % er.max())
        assert er.max() < 1e-5
        n = dev.frame_setup(hs.frame_desc(n_cores=1))
        assert n == int(meta["camera_samples"])
        dev.frame_samples()
        dev.frame_film_clear()
        Lr = dev.frame_render(0, n, hs.maxerror, hs.mix, want_L=True)
        order = dev.frame_order(0, n)
        L = np.empty_like(Lr); L[order] = Lr
        st = int(z["L_stride"])
        xy, _, _ = dev.frame_get_samples(0, n)
        assert np.array_equal(xy[::st].view(np.uint32), z["camera_xy"].view(np.uint32))
        got, ref = L[::st], z["L"]
        lit = ref.max(axis=1) > 0
        assert np.array_equal(got.max(axis=1) > 0, lit)
        e = gu.row_rel_err(got, ref)[lit]
        print("full Li with the transmission lobe: %d lit samples, within 1e-4: %.4f, within 1e-2: %.4f, max %.3g" % (lit.sum(), np.mean(e < 1e-4), np.mean(e < 1e-2), e.max()))
        assert e.max() < 1e-5, e.max()
        rgb = hs.film_to_rgb(dev.frame_film())
        _check_image(rgb, z["image"])
    finally:
        hs.close()


def _check_image(rgb, ref):
    assert rgb.shape == ref.shape
    for c in range(3):
        rmse = np.sqrt(np.mean((rgb[..., c] - ref[..., c]) ** 2)); scale = np.sqrt(np.mean(ref[..., c] ** 2))
        print("channel %d: RMSE / RMS = %.3g" % (c, rmse / scale))
        assert rmse < 1e-3 * scale


def test_tissue_pbrt_render(pkg, slab, golden):
    """The same frame through `pbrt scenes/tissue.pbrt` (parser + RenderScene in the C++ host)."""
    z, meta = golden
    rgb, st = pkg.pbrt_render(os.path.join(ROOT, "scenes", "tissue.pbrt"), n_cores=1)
    assert st["points"] == int(meta["surface_points"]) and st["samples"] == int(meta["camera_samples"])
    _check_image(rgb, z["image"])
