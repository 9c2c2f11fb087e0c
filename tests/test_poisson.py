"""MultipoleSubsurfaceIntegrator's "bool usepoissonpointfinder" (integrators/multipolesubsurface.cpp:183-185): the surface points come
from FindPoissonPointDistribution (renderers/surfacepoints.cpp:103-284,354-359) -- random walks through the scene that deposit
candidates, accepted when no accepted point lies within minSampleDist -- instead of the tessellator.  scenes/skin_poisson.scene
against the real reference's export with ONE SurfacePointTask (tests/golden/skin_poisson.npz, oracle/_ref/ref_driver --ncores 1;
with more tasks the reference's own result depends on thread timing).

The walks run on the device (k_poisson_walk: the scene's closest hit, the emitters' and the bounding sphere's Sphere::Intersect,
UniformSampleSphere through glibc's sinf / cosf restated), the serial accept / reject chain on the host (csrc/skc_poisson.cu)."""
import os

import numpy as np
import pytest

import golden_util as gu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def golden():
    z = np.load(os.path.join(ROOT, "tests", "golden", "skin_poisson.npz"))
    return z, dict(zip([str(k) for k in z["meta_keys"]], z["meta_vals"]))


def test_scene_formats_agree(pkg):
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import mesh_bin_to_pbrt
    dst = os.path.join(ROOT, "scenes", "geometry", "head.pbrt")
    if not os.path.exists(dst):
        mesh_bin_to_pbrt.convert(os.path.join(ROOT, "scenes", "head.mesh.bin"), dst)
    a = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_poisson.scene")); b = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_poisson.pbrt"))
    c = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_ld4.scene"))
    try:
        assert a.use_poisson and b.use_poisson and not c.use_poisson
        assert a.poisson_max_fails == b.poisson_max_fails == 2000
        assert np.array_equal(a.camera_position, b.camera_position)
        assert np.allclose(a.camera_position, [0.491171, -4.4848, 0.897406], atol=1e-6)
    finally:
        a.close(); b.close(); c.close()


@pytest.mark.gpu
def test_poisson_points_irradiance_and_frame(pkg, golden):
    z, meta = golden
    dev = pkg.SkinDevice(0)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_poisson.scene"))
    try:
        hs.upload(dev)
        hs.build_profile(dev)
        n_pts, stats = dev.poisson_resident(hs.camera_position, hs.minsampledistance, material_id=int(meta["skin_material_id"]), max_fails=hs.poisson_max_fails)
        ref = z["surface_points"]
        print("Poisson finder: %d points (reference %d), %d paths, %d rays, %d rounds" % (n_pts, len(ref), stats["paths"], stats["rays"], stats["rounds"]))
        assert n_pts == len(ref) == int(meta["surface_points"])
        pts = dev.get_surface_points(0, n_pts)
        got = np.frombuffer(pts.tobytes(), np.uint32).reshape(n_pts, 11)
        exact = np.all(got == ref.view(np.uint32), axis=1)
        print("points bit-exact (all 44 bytes): %.4f" % exact.mean())
        assert exact.all(), np.flatnonzero(~exact)[:5]
        dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
        _, E = dev.irradiance_rows(0, n_pts, 1)
        refE = z["irradiance_E"]
        e = gu.row_rel_err(E, refE)
        lit = refE.max(axis=1) > 0
        print("irradiance at the Poisson points: %d lit, bit-exact rows %.4f, max rel %.3g" % (lit.sum(), np.mean(e == 0), e.max()))
        assert np.array_equal(E.max(axis=1) > 0, lit) and e.max() < 1e-6
        dev.build_octree()
        n = dev.frame_setup(hs.frame_desc(n_cores=1))
        dev.frame_samples()
        dev.frame_film_clear()
        Lr = dev.frame_render(0, n, hs.maxerror, hs.mix, want_L=True)
        order = dev.frame_order(0, n)
        L = np.empty_like(Lr); L[order] = Lr
        refL = z["L"]
        lit = refL.max(axis=1) > 0
        assert np.array_equal(L.max(axis=1) > 0, lit)
        e = gu.row_rel_err(L, refL)[lit]
        print("full Li over the Poisson point set: %d lit samples, max rel %.3g" % (lit.sum(), e.max()))
        assert e.max() < 1e-4, e.max()
        rgb = hs.film_to_rgb(dev.frame_film())
        for c in range(3):
            rmse = np.sqrt(np.mean((rgb[..., c] - z["image"][..., c]) ** 2)); s = np.sqrt(np.mean(z["image"][..., c] ** 2))
            assert rmse < 1e-3 * s
    finally:
        hs.close()
        dev.close()


@pytest.mark.gpu
def test_surfacepoints_renderer_and_pbrt_executable(pkg, golden, tmp_path):
    """`Renderer "surfacepoints" "string filename"` writes the text file of SurfacePointsRenderer::Render (surfacepoints.cpp:150-166:
    two comment lines, then position, normal, area, rayEpsilon with %g) -- the golden's points printed the same way -- and
    `pbrt scenes/skin_poisson.pbrt` renders the reference's image from the finder's points."""
    import subprocess
    z, meta = golden
    exe = os.path.join(ROOT, "pbrt-v2-skin_b200", "pbrt")
    src = open(os.path.join(ROOT, "scenes", "skin_poisson.pbrt")).read()
    src = src.replace('Include "geometry/head.pbrt"', 'Include "%s"' % os.path.join(ROOT, "scenes", "geometry", "head.pbrt"))
    a = src.replace('SurfaceIntegrator "multipolesubsurface"', 'Renderer "surfacepoints" "float minsampledistance" 0.1 "string filename" "pts.txt"\nSurfaceIntegrator "multipolesubsurface"')
    (tmp_path / "points.pbrt").write_text(a)
    r = subprocess.run([exe, "--ncores", "1", str(tmp_path / "points.pbrt")], cwd=str(tmp_path), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    ref = z["surface_points"]
    assert "Obtained %d surface points" % len(ref) in r.stdout
    lines = open(tmp_path / "pts.txt").read().splitlines()
    assert lines[0] == "# points generated by SurfacePointsRenderer" and lines[1] == "# position (x,y,z), normal (x,y,z), area, rayEpsilon"
    want = ["%g %g %g %g %g %g %g %g" % tuple(float(v) for v in np.r_[p[:6], p[9:11]]) for p in ref]
    assert lines[2:] == want
    img, st = pkg.pbrt_render(os.path.join(ROOT, "scenes", "skin_poisson.pbrt"), n_cores=1)
    assert st["points"] == len(ref)
    for c in range(3):
        rmse = np.sqrt(np.mean((img[..., c] - z["image"][..., c]) ** 2)); s = np.sqrt(np.mean(z["image"][..., c] ** 2))
        assert rmse < 1e-3 * s
