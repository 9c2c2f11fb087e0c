"""The head scenes bind their textures as .tga files (results/PigmentManips/Eum0.1_Deoxy0.5.pbrt: lambertian-norm2.tga,
bump-lowRes2.tga; not shipped with the fork).  scenes/skin_tga.{scene,pbrt} bind synthetic stand-ins in that format
whose sizes are not powers of two (tools/make_tga_textures.py: a run-length encoded true-colour file with a top-left origin, a raw
grey file with a bottom-left origin), and the host mirror's ImageTexture -- ReadImageTGA (core/imageio.cpp:214-256), convertIn,
MIPMap's Lanczos zoom to the next power of two (core/mipmap.h:67-87,119-170) and the box-filtered levels -- has to equal the real
reference's pyramids bit for bit (tests/golden/tex_tga.npz, oracle/_ref/ref_driver).  CPU only: the device takes the pyramids as
they are (tests/test_albedo_texture.py, tests/test_bump.py pin that path)."""
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _split(flat, channels, has_lut):
    n = int(flat[:1].view(np.int32)[0]); o = 1
    w, h, levels = [], [], []
    for _ in range(n):
        lw, lh = [int(x) for x in flat[o:o + 2].view(np.int32)]; o += 2
        w.append(lw); h.append(lh); levels.append(flat[o:o + channels * lw * lh]); o += channels * lw * lh
    if has_lut:
        o += 128
    assert o == len(flat)
    return w, h, np.concatenate(levels)


@pytest.mark.parametrize("ext", ["scene", "pbrt", "b"])
def test_tga_textures_zoomed_pyramids_bit_exact(pkg, ext):
    """"b": the same pictures (the grey one with an extra flat patch) in the other encodings -- 32-bit raw stored right to left with a
    bottom-left origin, grey run-length encoded with a top-left origin (scenes/skin_tga_b.scene)."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "tex_tga.npz"))
    suffix = "_b" if ext == "b" else ""
    if ext == "pbrt":
        import sys
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import mesh_bin_to_pbrt
        dst = os.path.join(ROOT, "scenes", "geometry", "head.pbrt")
        if not os.path.exists(dst):
            mesh_bin_to_pbrt.convert(os.path.join(ROOT, "scenes", "head.mesh.bin"), dst)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_tga_b.scene" if ext == "b" else "skin_tga." + ext))
    try:
        a = hs.albedo_texture(); b = hs.bump_texture()
        w, h, rgb = _split(z["albedo_tex" + suffix], 3, True)
        assert list(a["w"]) == w == [64, 32, 16, 8, 4, 2, 1] and list(a["h"]) == h          # 48 x 40 zoomed to 64 x 64
        same = a["rgb"].view(np.uint32) == rgb.view(np.uint32)
        assert same.all(), "albedo pyramid: %d of %d floats differ, first at %d" % ((~same).sum(), same.size, int(np.argmin(same)))
        w, h, disp = _split(z["bump_tex" + suffix], 1, False)
        assert list(b["w"]) == w == [64, 32, 16, 8, 4, 2, 1] and list(b["h"]) == h
        same = b["texels"].view(np.uint32) == disp.view(np.uint32)
        assert same.all(), "bump pyramid: %d of %d floats differ, first at %d" % ((~same).sum(), same.size, int(np.argmin(same)))
    finally:
        hs.close()


def test_tga_reader_refuses_what_it_does_not_read(pkg, tmp_path):
    import struct
    bad = tmp_path / "cmap.tga"
    bad.write_bytes(struct.pack("<BBBHHBHHHHBB", 0, 1, 1, 0, 2, 24, 0, 0, 2, 2, 8, 0) + bytes(6) + bytes(4))      # colour-mapped
    scene = open(os.path.join(ROOT, "scenes", "skin_tga.scene")).read().replace("albedo48x40.tga", str(bad)).replace("head.mesh.bin", os.path.join(ROOT, "scenes", "head.mesh.bin")).replace("bump48x40.tga", os.path.join(ROOT, "scenes", "bump48x40.tga"))
    p = tmp_path / "bad.scene"
    p.write_text(scene)
    with pytest.raises(Exception) as e:
        pkg.HostScene(str(p))
    assert "colour-mapped" in str(e.value)
