"""LayeredSkin's "texture bumpmap" as a float image map (materials/layeredskin.cpp:144-145,246; core/material.cpp:47-104, this fork's
central-difference Material::Bump; shapes/trianglemesh.cpp:236-242): scenes/skin_bump.scene (the head with scenes/bump64.pfm,
tools/make_bumpmap.py) against the real reference's export (tests/golden/skin_bump.npz, oracle/_ref/ref_driver --ncores 1).

CPU: the host mirror's ImageTexture<float, float> with its MIPMap<float> pyramid equals the reference's bit for bit, from both scene
formats.  GPU: the bumped normal of every tessellated surface point (k_tess_emit: zero differentials, du = dv = .01), E of every
irradiance point (the normal decides the hemisphere and the cosine), the bumped shading normal of every camera hit (k_hit_textures: EWA
lookups under the ray's differentials) and the complete Li."""
import os

import numpy as np
import pytest

import golden_util as gu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def golden():
    z = np.load(os.path.join(ROOT, "tests", "golden", "skin_bump.npz"))
    return z, dict(zip([str(k) for k in z["meta_keys"]], z["meta_vals"]))


def _split_tex(flat):
    n = int(flat[:1].view(np.int32)[0]); o = 1
    w, h, levels = [], [], []
    for _ in range(n):
        lw, lh = [int(x) for x in flat[o:o + 2].view(np.int32)]; o += 2
        w.append(lw); h.append(lh); levels.append(flat[o:o + lw * lh]); o += lw * lh
    assert o == len(flat)
    return w, h, np.concatenate(levels)


@pytest.mark.parametrize("ext", ["scene", "pbrt"])
def test_bump_texture_construction_bit_exact(pkg, golden, ext):
    z, _ = golden
    if ext == "pbrt":
        import sys
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import mesh_bin_to_pbrt
        dst = os.path.join(ROOT, "scenes", "geometry", "head.pbrt")
        if not os.path.exists(dst):
            mesh_bin_to_pbrt.convert(os.path.join(ROOT, "scenes", "head.mesh.bin"), dst)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_bump." + ext))
    try:
        t = hs.bump_texture()
        w, h, texels = _split_tex(z["bump_tex"])
        assert list(t["w"]) == w == [64, 32, 16, 8, 4, 2, 1] and list(t["h"]) == h
        assert np.array_equal(t["texels"].view(np.uint32), texels.view(np.uint32)), "pyramid texels"
        assert t["map"] == (1.0, 1.0, 0.0, 0.0) and hs.albedo_texture() is None
    finally:
        hs.close()


@pytest.mark.gpu
def test_points_irradiance_and_frame_with_a_bump_map(pkg, golden):
    z, meta = golden
    dev = pkg.SkinDevice(0)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_bump.scene"))
    try:
        hs.upload(dev)
        hs.build_profile(dev)
        n_pts = dev.tessellate_resident(hs.minsampledistance, False, material_id=int(meta["skin_material_id"]))
        assert n_pts == int(meta["surface_points"])
        nrm = np.ascontiguousarray(dev.get_surface_points(0, n_pts)["n"])
        ref_n = z["surface_normals"]
        exact = np.mean(np.all(nrm.view(np.uint32) == ref_n.view(np.uint32), axis=1))
        dn = np.abs(nrm - ref_n).max()
        print("bumped surface-point normals: %d points, bit-exact %.4f, max abs diff %.3g" % (n_pts, exact, dn))
        assert dn < 1e-5, dn
        # the map does bend them: unbumped, a tenth of the normals would be off by more than 0.05
        dev.upload_bump_texture(None)
        dev.tessellate_resident(hs.minsampledistance, False, material_id=int(meta["skin_material_id"]))
        flat = dev.get_surface_points(0, n_pts)["n"]
        assert np.mean(np.abs(flat - ref_n).max(axis=1) > 0.05) > 0.1
        dev.upload_bump_texture(hs.bump_texture())
        dev.tessellate_resident(hs.minsampledistance, False, material_id=int(meta["skin_material_id"]))
        dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
        _, E = dev.irradiance_rows(0, n_pts, 1)
        ref = z["irradiance_E"]
        e = gu.row_rel_err(E, ref)
        lit = ref.max(axis=1) > 0
        print("irradiance at the bumped points: %d points (%d lit), bit-exact rows %.4f, max rel %.3g" % (len(ref), lit.sum(), np.mean(e == 0), e.max()))
        assert np.array_equal(E.max(axis=1) > 0, lit)
        assert e.max() < 1e-4, e.max()
        dev.build_octree()
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
        print("full Li with bumped BSDF frames: %d lit samples, max rel %.3g, within 1e-5: %.4f" % (lit.sum(), e.max(), np.mean(e < 1e-5)))
        assert e.max() < 1e-4, e.max()
        rgb = hs.film_to_rgb(dev.frame_film())
        for c in range(3):
            rmse = np.sqrt(np.mean((rgb[..., c] - z["image"][..., c]) ** 2)); s = np.sqrt(np.mean(z["image"][..., c] ** 2))
            print("channel %d: RMSE / RMS = %.3g" % (c, rmse / s))
            assert rmse < 1e-3 * s
    finally:
        hs.close()
        dev.close()
