"""The pbrt scene-description front end (pbrt-v2-skin_b200/host/skin_pbrt.cpp) -- boundary rows B1/B2/B3 of SURVEY.md 8b:
`pbrt scene.pbrt` with the reference's own scene language (core/pbrtlex.ll, core/pbrtparse.yy, core/api.cpp and the
Create* plugin factories' parameter names and defaults).

CPU tests: the parsed scene is compared, byte for byte, with the same scene given in the line format that the real
reference was driven with when the goldens were made (oracle/refbuild/ref_driver.cpp), so every parity result of the
.scene files carries over to the .pbrt files; plus the lexer / grammar / state-machine cases of the reference's
language and the refusals (anything the device path cannot render is an error that names the directive)."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


@pytest.fixture(scope="module")
def head_pbrt():
    """scenes/geometry/head.pbrt: the head mesh as a pbrt text block (a build product)."""
    import mesh_bin_to_pbrt
    dst = os.path.join(ROOT, "scenes", "geometry", "head.pbrt")
    if not os.path.exists(dst):
        mesh_bin_to_pbrt.convert(os.path.join(ROOT, "scenes", "head.mesh.bin"), dst)
    return dst


def bytes_of(a):
    return np.ascontiguousarray(a).view(np.uint8)


@pytest.mark.parametrize("name", ["skin_ld4", "skin"])
def test_pbrt_scene_equals_line_format_scene(pkg, head_pbrt, name):
    a = pkg.HostScene(os.path.join(ROOT, "scenes", name + ".scene"))
    b = pkg.HostScene(os.path.join(ROOT, "scenes", name + ".pbrt"))
    for k in ("nodes", "prims", "lights", "P", "N", "uv", "idx", "o2w", "skin"):
        assert np.array_equal(bytes_of(getattr(a, k)), bytes_of(getattr(b, k))), k
    assert (a.maxerror, a.minsampledistance, a.mix, a.maxdepth) == (b.maxerror, b.minsampledistance, b.mix, b.maxdepth)
    assert (a.xres, a.yres, a.spp, a.reverse, a.swaps) == (b.xres, b.yres, b.spp, b.reverse, b.swaps)
    for cores in (1, 16):       # camera matrices, filter table, task counts, Kr / Kt spectra
        assert np.array_equal(a.frame_desc(cores), b.frame_desc(cores))
    assert b.film_filename == name + ".pfm" and b.warnings == []


TRI = textwrap.dedent('''
    Material "layeredskin" "skinlayer layers" [0.25e6 1.4 20e6 1.4] %s
    Shape "trianglemesh" "point P" [0 0 0  1 0 0  0 1 0  1 1 0] "integer indices" [0 1 2 2 1 3] %s
''')
LIGHT = 'AttributeBegin AreaLightSource "area" "color L" [10 10 10] "integer nsamples" [8] Translate 0 0 3 Shape "sphere" "float radius" .25 AttributeEnd\n'


def scene_text(options='SurfaceIntegrator "multipolesubsurface"', world=None, mat="", shp=""):
    return options + "\nWorldBegin\n" + (LIGHT + TRI % (mat, shp) if world is None else world) + "\nWorldEnd\n"


def load(pkg, tmp_path, text, **kw):
    p = tmp_path / "t.pbrt"
    p.write_text(text)
    return pkg.HostScene(str(p), **kw)


def test_defaults_follow_the_reference_factories(pkg, tmp_path):
    s = load(pkg, tmp_path, scene_text())
    assert (s.xres, s.yres, s.spp) == (640, 480, 4)                       # film/image.cpp:240-241, lowdiscrepancy.cpp:87
    assert (s.maxdepth, s.maxerror, s.minsampledistance, s.mix) == (5, np.float32(.05), .25, .5)   # multipolesubsurface.cpp:309-316
    assert s.accelerator() == ("sah", 4)                                  # accelerators/bvh.cpp:549-550
    k = s.skin      # roughness nmperunit f_mel f_eu f_blood f_ohg desiredlength lerp (layeredskin.cpp:234-257)
    assert np.allclose(k[:8], [0.4, 100e6, 0.15, 1.0, 0.002, 0.3, 512, 1])     # roughness 0.4: layeredskin.cpp:234
    xw, yw, table = s.pixel_filter()
    assert (xw, yw) == (.5, .5) and np.all(table == 1)                    # box, api.cpp default
    assert list(s.light_samples()) == [8]
    assert s.n_tris == 2 and s.n_lights == 1 and s.film_filename == ""
    fd = s.frame_desc(1)
    kt = fd[4 * (5 + 16 + 16 + 2 + 256 + 90 + 1 + 2 + 30):][:120].view(np.float32)
    assert np.all(kt > 0.9), "Kt defaults to Spectrum(1) (layeredskin.cpp:236): the transmission lobe is on"


def test_lexer_and_grammar_forms(pkg, tmp_path):
    """single values without brackets, comments, rgb = color, string escapes, numbers like `.5`, `-1e-2`, `+3`, Include,
    Transform stacks and named coordinate systems."""
    inc = tmp_path / "sub"; inc.mkdir()
    (inc / "tri.pbrt").write_text(TRI % ("", '"float st" [0 0 1 0 0 1 1 1]'))
    opts = textwrap.dedent('''
        # a comment line
        Film "image" "integer xresolution" 32 "integer yresolution" [ 16 ] "string filename" "a\\\\b \\"q\\".pfm"  # trailing comment
        Scale -1 1 1 LookAt 0 0 5  0 0 0  0 1 0
        Camera "perspective" "float fov" 40 "float screenwindow" [-1 1 -.5 +.5]
        Sampler "lowdiscrepancy" "integer pixelsamples" 3
        PixelFilter "gaussian" "float xwidth" 1.5 "float alpha" 1
        Accelerator "bvh" "string splitmethod" "middle" "integer maxnodeprims" 2
        SurfaceIntegrator "multipolesubsurface" "float maxerror" 1e-1 "float mix" .25 "bool incenter" "true"
    ''')
    world = textwrap.dedent('''
        TransformBegin Translate 1 2 3 CoordinateSystem "there" TransformEnd
        AttributeBegin
          AreaLightSource "area" "rgb L" [4 4 4]
          CoordSysTransform "there"
          Shape "sphere"
        AttributeEnd
        AttributeBegin ReverseOrientation ConcatTransform [2 0 0 0  0 2 0 0  0 0 2 0  0 0 0 1]
          Include "sub/tri.pbrt"
        AttributeEnd
    ''')
    s = load(pkg, tmp_path, scene_text(opts, world))
    assert (s.xres, s.yres, s.spp) == (32, 16, 3) and s.film_filename == 'a\\b "q".pfm'
    assert (s.maxerror, s.mix) == (np.float32(.1), .25) and s.accelerator() == ("middle", 2)
    L = s.lights[0]
    assert tuple(L[[3, 7, 11]]) == (1, 2, 3), "the sphere sits at the named coordinate system's origin"
    assert s.reverse == 1 and s.uv is not None and np.array_equal(s.P[3], [2, 2, 0])
    xw, yw, t = s.pixel_filter()
    assert (xw, yw) == (1.5, 2.0)
    # GaussianFilter::Evaluate at the table's sample positions (filters/gaussian.cpp), fp32
    f = np.float32
    fx = (np.arange(16, dtype=f) + f(.5)) * f(1.5) / f(16); fy = (np.arange(16, dtype=f) + f(.5)) * f(2) / f(16)
    g = lambda d, w: np.maximum(f(0), np.exp(-f(1) * d * d, dtype=f) - np.exp(-f(1) * f(w) * f(w), dtype=f))
    assert np.allclose(t, np.outer(g(fy, 2), g(fx, 1.5)), rtol=1e-6, atol=1e-7)
    # screen window and the mirrored LookAt reach the camera matrices
    fd = s.frame_desc(1)
    c2w = fd[20:20 + 64].view(np.float32).reshape(4, 4)
    assert np.allclose(c2w[:3, 3], [0, 0, 5], atol=1e-6) and c2w[0, 0] == 1      # LookAt alone gives -1 (left = up x dir)


def test_quick_render_and_outfile_options(pkg, tmp_path):
    opts = 'Film "image" "integer xresolution" 64 "integer yresolution" 40\nSampler "lowdiscrepancy" "integer pixelsamples" 16\nSurfaceIntegrator "multipolesubsurface"'
    s = load(pkg, tmp_path, scene_text(opts), quick=True, outfile="cmdline.pfm")
    assert (s.xres, s.yres, s.spp) == (16, 10, 1)                         # film/image.cpp:242-243, lowdiscrepancy.cpp:88
    assert list(s.light_samples()) == [2]                                 # lights/diffuse.cpp:64
    assert s.film_filename == "cmdline.pfm"
    opts2 = opts.replace('"integer yresolution" 40', '"integer yresolution" 40 "string filename" "scene.tga"')
    s = load(pkg, tmp_path, scene_text(opts2), outfile="cmdline.pfm")
    assert s.film_filename == "scene.tga" and any("ignored" in w for w in s.warnings)   # film/image.cpp:222-231


def test_shape_parameters_override_material_parameters(pkg, tmp_path):
    s = load(pkg, tmp_path, scene_text(mat='"float f_mel" 0.3 "float roughness" 0.2', shp='"float f_mel" 0.01'))
    assert np.allclose(s.skin[[0, 2]], [0.2, 0.01])
    world = 'MakeNamedMaterial "face" "string type" "layeredskin" "skinlayer layers" [1e5 1.3 2e7 1.45] "float f_blood" 0.1\n' + LIGHT + \
            'NamedMaterial "face" Shape "trianglemesh" "point P" [0 0 0 1 0 0 0 1 0] "integer indices" [0 1 2]'
    s = load(pkg, tmp_path, scene_text(world=world))
    assert np.allclose(s.skin[[4, 8, 9, 10, 11]], [0.1, 1e5, 1.3, 2e7, 1.45])


@pytest.mark.parametrize("text,needle", [
    (scene_text('SurfaceIntegrator "path"'), 'SurfaceIntegrator "path"'),
    (scene_text(""), "multipolesubsurface"),
    (scene_text(world='LightSource "goniometric"\n' + TRI % ("", "")), 'LightSource "goniometric"'),
    (scene_text(mat='"texture albedo" "tex"'), "texture albedo"),
    (scene_text(mat='"color albedo" [0.8 0.6]'), "albedo"),      # a constant albedo is supported (tests/test_gpu_albedo.py); a malformed one is not
    (scene_text(mat='"bool doublerefsslf" "true"'), "doublerefsslf"),
    (scene_text(mat='"bool usemontecarlo" "true"'), "usemontecarlo"),
    (scene_text(mat='"bool genprofile" "false"'), "genprofile"),
    (scene_text('SurfaceIntegrator "multipolesubsurface" "bool showirradiancepoints" "true"'), "showirradiancepoints"),
    (scene_text('Renderer "metropolis"\nSurfaceIntegrator "multipolesubsurface"'), 'Renderer "metropolis"'),
    (scene_text(world=LIGHT + 'Shape "trianglemesh" "point P" [0 0 0 1 0 0 0 1 0] "integer indices" [0 1 2]'), 'Material "matte"'),
    (scene_text(world=LIGHT + 'Material "layeredskin"\nShape "trianglemesh" "point P" [0 0 0 1 0 0 0 1 0] "integer indices" [0 1 2]'), "No layers"),
    (scene_text(world=LIGHT + TRI % ("", "") + 'Shape "cylinder"'), 'Shape "cylinder"'),
    (scene_text(shp='"integer bogus [1]'), "Unterminated string"),
    (scene_text(world=LIGHT + (TRI % ("", "")).replace("[0 1 2 2 1 3]", "[0 1 2 2 1 4]")), "out of-bounds vertex index"),
    (scene_text('Camera "orthographic"\nSurfaceIntegrator "multipolesubsurface"'), 'Camera "orthographic"'),
    (scene_text('Sampler "random"\nSurfaceIntegrator "multipolesubsurface"'), 'Sampler "random"'),
    (scene_text(world='Include "nowhere.pbrt"'), "nowhere.pbrt"),
    ('SurfaceIntegrator "multipolesubsurface"\nShape "sphere"\n', "must be inside world block"),
    (scene_text(world=LIGHT + TRI % ("", "") + "\nFrobnicate 1 2"), "Frobnicate"),
    ('SurfaceIntegrator "multipolesubsurface"\nWorldBegin\n' + LIGHT + TRI % ("", ""), "WorldEnd"),
    (scene_text(shp='"floot x" 1'), "Unable to decode type"),
    (scene_text(world=LIGHT + 'ObjectBegin "o" ObjectEnd'), "instancing"),
])
def test_refusals_name_the_directive(pkg, tmp_path, text, needle):
    with pytest.raises(pkg.SkcError) as e:
        load(pkg, tmp_path, text)
    assert needle in str(e.value), str(e.value)
    assert "t.pbrt(" in str(e.value) or "scene" in str(e.value)


def test_unused_parameters_are_reported(pkg, tmp_path):
    s = load(pkg, tmp_path, scene_text('SurfaceIntegrator "multipolesubsurface" "float maxeror" 0.2'))
    assert any('"maxeror" not used' in w for w in s.warnings)             # ParamSet::ReportUnused
    s = load(pkg, tmp_path, scene_text(mat='"float rougness" 0.2 "float ga_epi" 0.9 "string photons" "100"'))
    assert any('"rougness" not used (Material' in w for w in s.warnings)  # the Material's own parameter list is checked too
    assert not any("ga_epi" in w or "photons" in w for w in s.warnings)    # read by CreateLayeredSkinMaterial (layeredskin.cpp:240-249)


def test_pbrt_executable_usage_and_errors(pkg, tmp_path):
    exe = os.path.join(ROOT, "pbrt-v2-skin_b200", "pbrt")
    assert os.path.exists(exe), "make -C pbrt-v2-skin_b200 builds the drop-in executable"
    out = subprocess.run([exe, "--help"], stdout=subprocess.PIPE, text=True)
    assert out.returncode == 0 and out.stdout.startswith("usage: pbrt [--ncores n] [--outfile filename] [--quick] [--quiet]")


@pytest.mark.gpu
def test_pbrt_executable_renders_the_reference_image(pkg, head_pbrt, tmp_path):
    """`pbrt --ncores 1 --quiet scenes/skin_ld4.pbrt` writes skin_ld4.pfm; against the real reference's image of the same
    scene (tests/golden/skin_ld4.npz): RMSE <= 1e-3 relative per channel (BASELINE.json)."""
    exe = os.path.join(ROOT, "pbrt-v2-skin_b200", "pbrt")
    r = subprocess.run([exe, "--ncores", "1", os.path.join(ROOT, "scenes", "skin_ld4.pbrt")], cwd=str(tmp_path), stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    print(r.stdout.strip())
    with open(tmp_path / "skin_ld4.pfm", "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = map(int, f.readline().split()); f.readline()
        img = np.frombuffer(f.read(), "<f4").reshape(h, w, 3)[::-1]
    ref = np.load(os.path.join(ROOT, "tests", "golden", "skin_ld4.npz"))["image"]
    assert img.shape == ref.shape
    for c in range(3):
        rmse = np.sqrt(np.mean((img[..., c] - ref[..., c]) ** 2)); scale = np.sqrt(np.mean(ref[..., c] ** 2))
        print("channel %d: RMSE / RMS = %.3g" % (c, rmse / scale))
        assert rmse < 1e-3 * scale
    # the library entry renders the same pixels, and a TGA comes out for a .tga name
    rgb, st = pkg.pbrt_render(os.path.join(ROOT, "scenes", "skin_ld4.pbrt"), n_cores=1)
    # (not bit for bit: the film accumulates with float atomics, whose order varies from run to run)
    assert np.allclose(rgb, img, rtol=1e-5, atol=1e-6 * img.max()) and st["samples"] == (w + 1) * (h + 1) * 4
    r = subprocess.run([exe, "--ncores", "1", "--quiet", "--quick", "--outfile", "q.tga", str(_no_filename(tmp_path))], cwd=str(tmp_path),
                       stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    tga = (tmp_path / "q.tga").read_bytes()
    assert tga[2] == 2 and tga[12] + 256 * tga[13] == 12 and tga[14] + 256 * tga[15] == 6 and len(tga) == 18 + 12 * 6 * 3


def _no_filename(tmp_path):
    src = open(os.path.join(ROOT, "scenes", "skin_ld4.pbrt")).read().replace(' "string filename" "skin_ld4.pfm"', "")
    src = src.replace('Include "geometry/head.pbrt"', 'Include "%s"' % os.path.join(ROOT, "scenes", "geometry", "head.pbrt"))
    p = tmp_path / "nofile.pbrt"
    p.write_text(src)
    return p


@pytest.mark.gpu
def test_tessellatesurfacepoints_renderer_and_pointsfile(pkg, head_pbrt, tmp_path):
    """`Renderer "tessellatesurfacepoints" "string filename"` writes the raw SurfacePoint[44 B] file the reference writes
    (renderers/surfacepoints.cpp:335-347) -- byte for byte the reference's own points (tests/golden/skin_tiny.npz, same mesh and
    minsampledistance) -- and `"string pointsfile"` renders from it (multipolesubsurface.cpp:176-181) to the same image."""
    exe = os.path.join(ROOT, "pbrt-v2-skin_b200", "pbrt")
    src = open(os.path.join(ROOT, "scenes", "skin_ld4.pbrt")).read()
    src = src.replace('Include "geometry/head.pbrt"', 'Include "%s"' % head_pbrt)
    a = src.replace('SurfaceIntegrator "multipolesubsurface"', 'Renderer "tessellatesurfacepoints" "float minsampledistance" 0.1 "string filename" "pts.bin"\nSurfaceIntegrator "multipolesubsurface"')
    (tmp_path / "points.pbrt").write_text(a)
    r = subprocess.run([exe, str(tmp_path / "points.pbrt")], cwd=str(tmp_path), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    raw = np.fromfile(tmp_path / "pts.bin", np.uint8)       # "filename" resolves beside the scene file (core/fileutil.cpp:61-73)
    ref = np.load(os.path.join(ROOT, "tests", "golden", "skin_tiny.npz"))["surface_points"]
    assert "Obtained %d surface points" % (len(ref) // 44) in r.stdout
    assert np.array_equal(raw, ref)
    b = src.replace('"float minsampledistance" 0.1', '"float minsampledistance" 0.1 "string pointsfile" "pts.bin"').replace("skin_ld4.pfm", "from_file.pfm")
    (tmp_path / "fromfile.pbrt").write_text(b)
    img_file, st = pkg.pbrt_render(str(tmp_path / "fromfile.pbrt"), n_cores=1)
    img_tess, _ = pkg.pbrt_render(os.path.join(ROOT, "scenes", "skin_ld4.pbrt"), n_cores=1)
    assert st["points"] == len(ref) // 44
    assert np.allclose(img_file, img_tess, rtol=1e-5, atol=1e-6 * img_tess.max())
    # a missing points file falls back to the tessellator, as pts.size() == 0 does in the reference (:182-190)
    (tmp_path / "missing.pbrt").write_text(b.replace('"pts.bin"', '"nothere.bin"'))
    img_missing, _ = pkg.pbrt_render(str(tmp_path / "missing.pbrt"), n_cores=1)
    assert np.allclose(img_missing, img_tess, rtol=1e-5, atol=1e-6 * img_tess.max())


def test_several_skin_shapes_join_into_the_device_mesh(pkg, tmp_path):
    """Two Shape "trianglemesh" directives under one layeredskin material (the slab's top and bottom quad given separately)
    build the same device scene as the one-mesh file: vertices, indices, primitive list and BVH byte for byte."""
    import make_slab_mesh
    n = 12
    P, idx = make_slab_mesh.slab(n)
    nv = (n + 1) ** 2
    fmt = lambda a: " ".join("%.9g" % float(v) for v in a.ravel())
    head = ('SurfaceIntegrator "multipolesubsurface" "float minsampledistance" 0.02\nWorldBegin\n' + LIGHT +
            'Material "layeredskin" "skinlayer layers" [0.25e6 1.4 20e6 1.4]\n')
    one = head + 'Shape "trianglemesh" "point P" [%s] "integer indices" [%s]\nWorldEnd\n' % (fmt(P), " ".join(map(str, idx.ravel())))
    top, bot = idx[:len(idx) // 2], idx[len(idx) // 2:] - nv
    two = head + 'Shape "trianglemesh" "point P" [%s] "integer indices" [%s]\n' % (fmt(P[:nv]), " ".join(map(str, top.ravel()))) + \
        'Shape "trianglemesh" "point P" [%s] "integer indices" [%s]\nWorldEnd\n' % (fmt(P[nv:]), " ".join(map(str, bot.ravel())))
    (tmp_path / "one.pbrt").write_text(one); (tmp_path / "two.pbrt").write_text(two)
    a, b = pkg.HostScene(str(tmp_path / "one.pbrt")), pkg.HostScene(str(tmp_path / "two.pbrt"))
    assert b.n_tris == 4 * n * n and b.n_verts == 2 * nv
    for k in ("nodes", "prims", "P", "idx", "lights"):
        assert np.array_equal(bytes_of(getattr(a, k)), bytes_of(getattr(b, k))), k
    # meshes that cannot share the device's one transform / orientation are refused
    bad = two.replace('Shape "trianglemesh" "point P" [%s]' % fmt(P[nv:]), 'ReverseOrientation\nShape "trianglemesh" "point P" [%s]' % fmt(P[nv:]))
    (tmp_path / "bad.pbrt").write_text(bad)
    with pytest.raises(pkg.SkcError) as e:
        pkg.HostScene(str(tmp_path / "bad.pbrt"))
    assert "cannot be joined" in str(e.value)
