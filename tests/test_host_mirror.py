"""CPU tests of the C++ host mirror against the reference exports (tests/golden/skin_tiny.npz):
the flat BVH, the world-space mesh, the light records, the tessellated surface points and the
camera rays must be BIT-IDENTICAL to what the reference builds for the same scene file."""
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host(pkg):
    s = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_tiny.scene"))
    yield s
    s.close()


def test_world_mesh_bit_exact(host, tiny):
    assert np.array_equal(host.P.view(np.uint32), tiny.mesh_P.view(np.uint32))
    assert np.array_equal(host.idx, tiny.mesh_idx)
    assert np.array_equal(host.o2w.view(np.uint32), tiny.mesh_o2w.view(np.uint32))
    assert host.reverse == int(tiny.meta["mesh0_reverse"]) and host.swaps == int(tiny.meta["mesh0_swaps"])


def test_lights_bit_exact(host, tiny):
    assert np.array_equal(host.lights.view(np.uint32), tiny.lights.view(np.uint32))


def test_bvh_identical(host, tiny):
    assert host.n_nodes == len(tiny.bvh_nodes)
    assert np.array_equal(host.prims[:, [0, 2]], tiny.bvh_prims[:, [0, 2]])
    a, b = host.nodes, tiny.bvh_nodes
    for f in ("pmin", "pmax"):
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f
    assert np.array_equal(a["offset"], b["offset"]) and np.array_equal(a["nPrimitives"], b["nPrimitives"])
    interior = b["nPrimitives"] == 0
    assert np.array_equal(a["axis"][interior], b["axis"][interior])


def test_surface_points_bit_exact(host, tiny):
    pts = host.tessellate()
    ref = tiny.surface_points
    assert len(pts) == len(ref) == host.count_points()
    for f in ("p", "area", "rayEpsilon", "u", "v"):
        assert np.array_equal(pts[f].view(np.uint32), ref[f].view(np.uint32)), f
    assert np.array_equal(pts["materialId"], ref["materialId"])
    assert np.abs(pts["n"] - ref["n"]).max() < 1e-6
    assert np.mean((pts["n"].view(np.uint32) == ref["n"].view(np.uint32)).all(1)) > 0.99


def test_camera_rays_bit_exact(host, tiny):
    rays = host.camera_rays(tiny.camera_xy)
    assert np.array_equal(rays.view(np.uint32), tiny.camera_rays.view(np.uint32))


def test_irradiance_tasks(host, tiny):
    assert host.irradiance_tasks(len(tiny.surface_points), 1) == tiny.n_tasks
    assert host.irradiance_tasks(24648458, 8) == 8192


def test_bench_rays_cover_film(host):
    rays = host.bench_rays()
    assert len(rays) == host.xres * host.yres * host.spp
    assert np.isfinite(rays[:, :7]).all()


def test_skin_layer_coefficients_bit_exact(pkg, tiny):
    """a11: pigment model -> 30-bin mua/musp per layer, eta, thickness (materials/layeredskin.cpp:49-92,
    materials/skincoeffs.h:46-135) against the coefficients the reference's LayeredSkin computed."""
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin_tiny.scene"))
    try:
        c = hs.skin_coefficients()
        assert np.array_equal(c["flat"], tiny.skin_coeffs)
        assert c["mua"].shape == (2, 30) and np.all(c["mua"] > 0) and np.all(c["musp"] > 0)
    finally:
        hs.close()
