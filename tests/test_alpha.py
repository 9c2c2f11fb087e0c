"""Alpha-tested triangles (shapes/trianglemesh.inl:117-126,174-207): the head with "texture alpha" = a float image
map with exact zeros (scenes/alpha32.pfm, tools/make_alphamap.py) -- scenes/skin_alpha.scene against the real reference's export
(tests/golden/skin_alpha.npz, oracle/_ref/ref_driver --ncores 1).  Every kind of ray passes through the holes: camera rays (hits
bit-exact), the irradiance pass's shadow rays (E rows), both rays of EstimateDirect (full Li), the sampler stream's any-hit walks.
Bars (BASELINE.json): BVH hits bit-exact, irradiance 1e-4 per point, image RMSE 1e-3 per channel."""
import os

import numpy as np
import pytest

import golden_util as gu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def golden():
    z = np.load(os.path.join(ROOT, "tests", "golden", "skin_alpha.npz"))
    return z, dict(zip([str(k) for k in z["meta_keys"]], z["meta_vals"]))


@pytest.mark.parametrize("ext", ["scene", "pbrt"])
def test_alpha_texture_construction_bit_exact(pkg, golden, ext):
    """ImageTexture<float, float> in the host mirror (PFM reader, luminance, convertIn) = the reference's MIPMap level 0."""
    z, _ = golden
    if ext == "pbrt":
        import sys
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import mesh_bin_to_pbrt
        dst = os.path.join(ROOT, "scenes", "geometry", "head_alpha.pbrt")
        if not os.path.exists(dst):
            mesh_bin_to_pbrt.convert(os.path.join(ROOT, "scenes", "head.mesh.bin"), dst, extra='"texture alpha" "alphamap"')
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_alpha." + ext))
    try:
        a = hs.alpha_texture()
        w, h = [int(x) for x in z["alpha"][:2].view(np.int32)]
        assert (a["w"], a["h"]) == (w, h) == (32, 32) and a["map"] == (1.0, 1.0, 0.0, 0.0)
        assert np.array_equal(a["texels"].ravel().view(np.uint32), z["alpha"][2:].view(np.uint32))
        assert (a["texels"] == 0).mean() > 0.1, "the mask has exact zeros"
    finally:
        hs.close()


@pytest.mark.gpu
def test_alpha_tested_head_vs_reference(pkg, golden):
    z, meta = golden
    dev = pkg.SkinDevice(0)         # its own context: the alpha texture stays on a context until it is replaced
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_alpha.scene"))
    try:
        hs.upload(dev)
        hs.build_profile(dev)
        # camera rays of the reference through the alpha-tested BVH: closest hits bit-exact, any-hit agrees
        rays = np.ascontiguousarray(z["camera_rays"])
        ref_hits = z["hits"].view(gu.DT_HIT).reshape(-1)
        got = dev.trace_closest(rays)
        m = ref_hits["hit"] == 1
        assert np.array_equal(got["hit"], ref_hits["hit"])
        assert np.array_equal(got["prim"][m], ref_hits["prim"][m])
        assert np.array_equal(got["t"][m].view(np.uint32), ref_hits["t"][m].view(np.uint32)), "t not bit-exact"
        sk = ref_hits["hasBssrdf"] == 1
        assert np.array_equal(got["p"][sk].view(np.uint32), ref_hits["p"][sk].view(np.uint32)), "hit point not bit-exact"
        assert np.array_equal(dev.trace_any(rays) != 0, ref_hits["hit"] != 0)
        # ... and they differ from the opaque head: some rays go through a hole
        dev.upload_alpha(None)
        opaque = dev.trace_closest(rays)
        through = (opaque["hit"] != 0) & ((got["hit"] == 0) | (got["prim"] != opaque["prim"]))
        print("camera rays that pass through a hole of the mask: %d of %d hits" % (through.sum(), (opaque["hit"] != 0).sum()))
        assert through.sum() > 50
        dev.upload_alpha(hs.alpha_texture())
        n_pts = dev.tessellate_resident(hs.minsampledistance, False, material_id=int(meta["skin_material_id"]))
        assert n_pts == int(meta["surface_points"])
        dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
        _, E = dev.irradiance_rows(0, n_pts, 1)
        ref = z["irradiance_E"]
        e = gu.row_rel_err(E, ref)
        lit = ref.max(axis=1) > 0
        print("irradiance behind the alpha mask: %d points (%d lit), bit-exact rows %.4f, max rel %.3g" % (len(ref), lit.sum(), np.mean(e == 0), e.max()))
        assert np.array_equal(E.max(axis=1) > 0, lit)
        assert e.max() < 1e-6, e.max()
        dev.build_octree()
        n = dev.frame_setup(hs.frame_desc(n_cores=1))
        assert n == int(meta["camera_samples"])
        dev.frame_samples()
        xy, _, _ = dev.frame_get_samples(0, n)
        assert np.array_equal(xy.view(np.uint32), z["camera_xy"].view(np.uint32)), "the sampler stream depends on which camera rays hit"
        dev.frame_film_clear()
        Lr = dev.frame_render(0, n, hs.maxerror, hs.mix, want_L=True)
        order = dev.frame_order(0, n)
        L = np.empty_like(Lr); L[order] = Lr
        ref = z["L"]
        lit = ref.max(axis=1) > 0
        assert np.array_equal(L.max(axis=1) > 0, lit)
        e = gu.row_rel_err(L, ref)[lit]
        print("full Li: %d lit samples, max rel %.3g" % (lit.sum(), e.max()))
        assert e.max() < 1e-5, e.max()
        rgb = hs.film_to_rgb(dev.frame_film())
        for c in range(3):
            rmse = np.sqrt(np.mean((rgb[..., c] - z["image"][..., c]) ** 2)); s = np.sqrt(np.mean(z["image"][..., c] ** 2))
            print("channel %d: RMSE / RMS = %.3g" % (c, rmse / s))
            assert rmse < 1e-3 * s
    finally:
        hs.close()
        dev.close()
