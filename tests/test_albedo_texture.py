"""LayeredSkin's "texture albedo" as an image map (materials/layeredskin.cpp:183,247; textures/imagemap.cpp; core/mipmap.h):
scenes/skin_tex.scene (the head with scenes/albedo64.pfm, tools/make_albedomap.py, mix 0.4) against the real reference's export
(tests/golden/skin_tex.npz, oracle/_ref/ref_driver --ncores 1).

CPU: the host mirror's ImageTexture<RGBSpectrum, Spectrum> (PFM reader, convertIn, the MIPMap pyramid, the EWA weight table) equals
the reference's bit for bit.  GPU: E of every irradiance point (albedo looked up with zero differentials, Spectrum::FromRGB, powf
restated on the device) and the complete Li (EWA lookup under the camera ray's differentials)."""
import os

import numpy as np
import pytest

import golden_util as gu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def golden():
    z = np.load(os.path.join(ROOT, "tests", "golden", "skin_tex.npz"))
    return z, dict(zip([str(k) for k in z["meta_keys"]], z["meta_vals"]))


def _split_tex(flat):
    n = int(flat[:1].view(np.int32)[0]); o = 1
    w, h, levels = [], [], []
    for _ in range(n):
        lw, lh = [int(x) for x in flat[o:o + 2].view(np.int32)]; o += 2
        w.append(lw); h.append(lh); levels.append(flat[o:o + 3 * lw * lh]); o += 3 * lw * lh
    lut = flat[o:o + 128]; o += 128
    assert o == len(flat)
    return w, h, np.concatenate(levels), lut


@pytest.mark.parametrize("ext", ["scene", "pbrt"])
def test_albedo_texture_construction_bit_exact(pkg, golden, ext):
    z, _ = golden
    if ext == "pbrt":
        import sys
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import mesh_bin_to_pbrt
        dst = os.path.join(ROOT, "scenes", "geometry", "head.pbrt")
        if not os.path.exists(dst):
            mesh_bin_to_pbrt.convert(os.path.join(ROOT, "scenes", "head.mesh.bin"), dst)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_tex." + ext))
    try:
        t = hs.albedo_texture()
        w, h, rgb, lut = _split_tex(z["albedo_tex"])
        assert list(t["w"]) == w == [64, 32, 16, 8, 4, 2, 1] and list(t["h"]) == h
        assert np.array_equal(t["rgb"].view(np.uint32), rgb.view(np.uint32)), "pyramid texels"
        assert np.array_equal(t["lut"].view(np.uint32), lut.view(np.uint32)), "EWA weight table"
        assert t["map"] == (1.0, 1.0, 0.0, 0.0) and hs.mix == np.float32(0.4)
    finally:
        hs.close()


@pytest.mark.gpu
def test_device_powf_equals_libm(dev):
    """Pow(albedo, mix) is powf per wavelength (core/spectrum.h); the device restates glibc's (csrc/skc_common.cuh skc_powf)."""
    import ctypes
    libm = ctypes.CDLL("libm.so.6")
    f = libm.powf; f.restype = ctypes.c_float; f.argtypes = [ctypes.c_float, ctypes.c_float]
    rng = np.random.default_rng(5)
    x = np.concatenate([rng.uniform(1e-3, 1.2, 100000), rng.random(20000) ** 6, [1.0, 0.5, 0.94, 1e-30, 2.0 ** -130]]).astype(np.float32)
    y = np.concatenate([rng.choice([0.3, 0.4, 0.5, 0.6, 0.7, 1.0], 60000), rng.random(60005)]).astype(np.float32)
    pairs = np.concatenate([x, y[::-1]]).astype(np.float32)          # kernel: powf(v[i], v[n - 1 - i])
    n = len(pairs)
    host = np.array([f(float(pairs[i]), float(pairs[n - 1 - i])) for i in range(len(x))], np.float32)
    got = dev.test_libmf(7, pairs)[:len(x)]
    assert np.array_equal(got.view(np.uint32), host.view(np.uint32))


@pytest.mark.gpu
def test_irradiance_and_frame_with_an_albedo_texture(pkg, golden):
    z, meta = golden
    dev = pkg.SkinDevice(0)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_tex.scene"))
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
        print("irradiance with a textured albedo: %d points (%d lit), bit-exact rows %.4f, max rel %.3g" % (len(ref), lit.sum(), np.mean(e == 0), e.max()))
        assert np.array_equal(E.max(axis=1) > 0, lit)
        assert e.max() < 1e-6, e.max()
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
        print("full Li with the EWA-filtered albedo: %d lit samples, max rel %.3g, within 1e-5: %.4f" % (lit.sum(), e.max(), np.mean(e < 1e-5)))
        assert e.max() < 1e-4, e.max()
        rgb = hs.film_to_rgb(dev.frame_film())
        for c in range(3):
            rmse = np.sqrt(np.mean((rgb[..., c] - z["image"][..., c]) ** 2)); s = np.sqrt(np.mean(z["image"][..., c] ** 2))
            print("channel %d: RMSE / RMS = %.3g" % (c, rmse / s))
            assert rmse < 1e-3 * s
    finally:
        hs.close()
        dev.close()
