"""BASELINE.json configs[3] at a size the CPU oracle finishes in seconds: the synthetic Mo workload (icosphere + analytic
irradiance field) and the BVH microbenchmark's ray sets, device against oracle/skin_oracle.c.
Bars: BVH hit flag / primitive / t bit-exact; Mo within 1e-4 relative (BASELINE.json); octree root bit-exact."""
import os
import sys
import types

import numpy as np
import pytest

import golden_util as gu
import oracle_api as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ico(pkg, dev, tmp_path_factory):
    import synthetic as syn
    d = str(tmp_path_factory.mktemp("ico"))
    hs = pkg.HostScene(syn.write_scene(d, 4, film=64, desired_length=64))
    assert hs.n_tris == 20 * 4 ** 4
    hs.upload(dev)
    hdr, rd = hs.build_profile(dev)
    g = types.SimpleNamespace(bvh_nodes=hs.nodes, bvh_prims=hs.prims, mesh_P=hs.P, mesh_N=hs.N, mesh_uv=hs.uv, mesh_idx=hs.idx, lights=hs.lights,
                              mesh_o2w=hs.o2w, meta={"mesh0_reverse": hs.reverse, "mesh0_swaps": hs.swaps})
    yield syn, hs, g, hdr, rd
    hs.close()


def test_icosphere_is_closed_and_outward(ico):
    syn, hs, g, _, _ = ico
    P, F = syn.icosphere(3)
    a, b, c = P[F[:, 0]], P[F[:, 1]], P[F[:, 2]]
    assert np.all(np.einsum("ij,ij->i", np.cross(b - a, c - a), a) > 0) and len(P) == 10 * 4 ** 3 + 2
    e = np.sort(np.concatenate([F[:, [0, 1]], F[:, [1, 2]], F[:, [2, 0]]]), axis=1)
    _, counts = np.unique(e, axis=0, return_counts=True)
    assert np.all(counts == 2), "every edge is shared by exactly two triangles"


@pytest.mark.parametrize("dist", ["aggregate", "through_mesh"])
def test_bvh_ray_sets_bit_exact_vs_oracle(dev, ico, dist):
    syn, hs, g, _, _ = ico
    n = 20000
    rays = syn.aggregate_test_rays(hs.P.min(axis=0), hs.P.max(axis=0), n) if dist == "aggregate" else syn.through_mesh_rays(hs.P, n)
    dev.set_camera_rays(rays)
    hits, c = dev.trace_resident(0, n, want_out=True, counters=True)
    ref, cref = orc.trace_closest(g, rays, counters=True)
    assert np.array_equal(hits["hit"], ref["hit"]) and np.array_equal(hits["prim"], ref["prim"])
    assert np.array_equal(hits["t"].view(np.uint32), ref["t"].view(np.uint32)) and np.array_equal(hits["p"].view(np.uint32), ref["p"].view(np.uint32))
    assert np.array_equal(c, cref), "the device visits exactly the nodes and primitives the reference's loop does"
    occ, _ = dev.trace_resident(0, n, any_hit=True, want_out=True)
    assert np.array_equal(occ != 0, orc.trace_any(g, rays) != 0)
    if dist == "through_mesh":
        assert hits["hit"].mean() > 0.95      # rays aimed at silhouette vertices can slip past
    print("%s: %.3f of the rays hit, %.1f slab tests and %.2f primitive tests per ray" % (dist, hits["hit"].mean(), c[0] / n, c[1] / n))


@pytest.mark.parametrize("max_error", [0.05, 0.2])
def test_synthetic_mo_vs_oracle(pkg, dev, ico, max_error):
    syn, hs, g, hdr, rd = ico
    n_pts = dev.tessellate_resident(0.012, False, material_id=2)
    assert 1e5 < n_pts < 1e6
    sp = dev.get_surface_points(0, n_pts)
    ip = syn.synthetic_irradiance(sp, pkg.DT_IRR_POINT)
    assert (ip["E"].max(axis=1) > 0).mean() > 0.45 and np.all(ip["E"][:, 29] >= ip["E"][:, 0])
    dev.set_irradiance_points(ip)
    n_nodes, root, bounds = dev.build_octree()
    cie_y = hs.frame_desc(1)[4 * (5 + 16 + 16 + 2 + 256 + 30):][:120].view(np.float32)
    tree = orc.Octree(ip.view(gu.DT_IP), cie_y)
    n_ref, root_ref, bounds_ref = tree.info()
    assert n_nodes == n_ref and np.array_equal(bounds.view(np.uint32), bounds_ref.view(np.uint32))
    assert np.array_equal(root.view(np.uint32), root_ref.view(np.uint32))
    pts = syn.random_hit_points(sp, 3000)
    got = dev.mo(pts, max_error)
    ref, cnt = tree.mo(pts, rd, hdr[:, 1], max_error, counters=True)
    e = gu.row_rel_err(got, ref)
    print("maxError %.2f: %d nodes, %.0f node visits / %.0f contributions per hit, max rel %.3g" % (max_error, n_nodes, cnt[0] / len(pts), (cnt[1] + cnt[2]) / len(pts), e.max()))
    assert e.max() < 1e-4
    tree.close()
