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
