"""The synthetic workload generators of BASELINE.json configs[3] / [4] (tools/synthetic.py, tools/profile_batch_bench.py,
tools/make_slab_mesh.py): deterministic, well-formed inputs.  CPU tests (no device)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import synthetic as syn  # noqa: E402
import make_slab_mesh  # noqa: E402


def test_icosphere_is_closed_outward_and_sized():
    for level in (0, 2, 4):
        P, F = syn.icosphere(level)
        assert len(F) == 20 * 4 ** level and len(P) == 10 * 4 ** level + 2
        assert np.allclose(np.linalg.norm(P, axis=1), 1, atol=1e-6)
        a, b, c = P[F[:, 0]], P[F[:, 1]], P[F[:, 2]]
        assert np.all(np.einsum("ij,ij->i", np.cross(b - a, c - a), a) > 0), "outward-facing"
        e = np.sort(np.concatenate([F[:, [0, 1]], F[:, [1, 2]], F[:, [2, 0]]]), axis=1)
        _, counts = np.unique(e, axis=0, return_counts=True)
        assert np.all(counts == 2), "closed: every edge shared by exactly two triangles"


def test_ray_sets_are_deterministic_and_normalised():
    bmin, bmax = np.array([-1, -1, -1], np.float32), np.array([1, 1, 1], np.float32)
    r1, r2 = syn.aggregate_test_rays(bmin, bmax, 5000), syn.aggregate_test_rays(bmin, bmax, 5000)
    assert np.array_equal(r1, r2)
    assert np.allclose(np.linalg.norm(r1[:, 3:6], axis=1), 1, atol=1e-6) and np.all(np.isinf(r1[:, 7]))
    axis = ((r1[:, 3:6] == 0).sum(1) == 2).mean()
    assert 0.07 < axis < 0.12, "3/32 of the directions are snapped to an axis (aggregatetest.cpp:79-82)"
    assert np.all(r1[:, :3] >= -3) and np.all(r1[:, :3] <= 3), "origins inside the bound grown by its largest extent"
    P, _ = syn.icosphere(3)
    t = syn.through_mesh_rays(P, 2000)
    o, d = t[:, :3].astype(np.float64), t[:, 3:6].astype(np.float64)
    b = (o * d).sum(1); disc = b * b - ((o * o).sum(1) - 1)
    assert np.allclose(np.linalg.norm(o, axis=1), 2, atol=1e-5) and (disc > -1e-6).mean() > 0.99, "origins on the 2x sphere, aimed at the mesh"


def test_slab_mesh_and_irradiance_field():
    P, idx = make_slab_mesh.slab(10)
    assert len(P) == 2 * 11 * 11 and len(idx) == 4 * 100 and idx.max() == len(P) - 1
    a, b, c = P[idx[:, 0]], P[idx[:, 1]], P[idx[:, 2]]
    nz = np.cross(b - a, c - a)[:, 2]
    assert np.all(nz[:200] > 0) and np.all(nz[200:] < 0), "top faces +z, bottom faces -z"
    dt_sp = np.dtype([("p", "<f4", 3), ("n", "<f4", 3), ("u", "<f4"), ("v", "<f4"), ("materialId", "<u4"), ("area", "<f4"), ("rayEpsilon", "<f4")])
    dt_ip = np.dtype([("p", "<f4", 3), ("n", "<f4", 3), ("E", "<f4", 30), ("area", "<f4"), ("rayEpsilon", "<f4")])
    sp = np.zeros(1000, dt_sp)
    sp["n"][:, 2] = np.linspace(-1, 1, 1000); sp["area"] = 1
    ip = syn.synthetic_irradiance(sp, dt_ip)
    assert np.all(ip["E"][sp["n"][:, 2] <= 0] == 0) and np.all(ip["E"][sp["n"][:, 2] > 0.01] > 0)
    assert np.allclose(ip["E"][:, 29], ip["E"][:, 0] * np.float32(1.1), rtol=1e-6)
    assert np.array_equal(ip["E"], syn.synthetic_irradiance(sp, dt_ip)["E"])


def test_c5_parameter_sets_are_thin(pkg):
    import profile_batch_bench
    specs, lps, opts = profile_batch_bench.c5_parameter_sets(pkg, n_thick=4, desired_length=64)
    assert len(lps) == 120 and len(specs) == 240 and all(l == 2 for l in lps)
    for s in range(0, 120, 7):
        epi = specs[2 * s]
        mfp = 1.0 / (epi[2] + epi[3])
        assert 0.019 * mfp <= epi[1] <= 2.001 * mfp, "epidermis between 0.02 and 2 mean free paths: the thin-slab branch"
        assert opts[s][1] == 64 and opts[s][2] == 1
