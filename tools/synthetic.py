"""Synthetic workloads of BASELINE.json configs[3] (SURVEY.md 8d, C4 + the BVH microbenchmark): icosphere meshes, the
analytic irradiance field, random hit-point and ray sets.  Workload generation only -- shared by tools/synthetic_sweep.py
(timing on the B200) and tests/test_gpu_synthetic.py (parity at a small size)."""
import os

import numpy as np


def icosphere(level):
    """Unit icosphere, 20 * 4**level triangles (level 8: 1 310 720), outward-facing, vertices shared."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    V = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], np.float64)
    V /= np.linalg.norm(V, axis=1, keepdims=True)
    F = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
                  [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], np.int64)
    for _ in range(level):
        e = np.concatenate([F[:, [0, 1]], F[:, [1, 2]], F[:, [2, 0]]])
        key = np.sort(e, axis=1)
        uniq, inv = np.unique(key[:, 0] * (len(V) + 1) + key[:, 1], return_inverse=True)
        a, b = uniq // (len(V) + 1), uniq % (len(V) + 1)
        mid = V[a] + V[b]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        m = inv.reshape(3, -1) + len(V)          # midpoint vertex of edges 01, 12, 20 per face
        V = np.concatenate([V, mid])
        F = np.concatenate([np.stack([F[:, 0], m[0], m[2]], 1), np.stack([F[:, 1], m[1], m[0]], 1),
                            np.stack([F[:, 2], m[2], m[1]], 1), np.stack([m[0], m[1], m[2]], 1)])
    return V.astype(np.float32), F.astype(np.int32)


def write_mesh_bin(path, P, idx, N=None):
    with open(path, "wb") as f:
        np.array([len(P), len(idx), 0 if N is None else 1, 0], np.int32).tofile(f)
        np.ascontiguousarray(P, np.float32).tofile(f)
        if N is not None:
            np.ascontiguousarray(N, np.float32).tofile(f)
        np.ascontiguousarray(idx, np.int32).tofile(f)


def write_scene(dirname, level, min_dist=0.01, film=2048, spp=1, with_normals=True, desired_length=512):
    """Unit icosphere + one layeredskin material at its defaults + two sphere emitters, perspective camera at distance 3,
    fov 40.  Returns the scene path."""
    P, idx = icosphere(level)
    os.makedirs(dirname, exist_ok=True)
    write_mesh_bin(os.path.join(dirname, "ico%d.mesh.bin" % level), P, idx, P if with_normals else None)
    path = os.path.join(dirname, "ico%d.scene" % level)
    with open(path, "w") as f:
        f.write("film %d %d\nlookat 0 -3 0 0 0 0 0 0 1\nfov 40\nspp %d\n" % (film, film, spp))
        f.write("integrator maxdepth 5 maxerror 0.1 minsampledistance %r\n" % min_dist)
        f.write("light 0.5 -4 4 0.2 100 100 100 4\nlight 3 4 1 0.2 100 100 100 4\n")
        f.write("skin roughness 0.3 nmperunit 100e6 f_mel 0.15 f_eu 1 f_blood 0.002 f_ohg 0.3 desiredlength %d Kr 1 1 1 Kt 1 1 1 layers 0.25e6 1.4 20e6 1.4\n" % desired_length)
        f.write("mesh ico%d.mesh.bin\n" % level)
    return path


def synthetic_irradiance(points, dt_irr):
    """E_lambda(i) = max(0, n_i . l) (0.5 + u_i) (1 + 0.1 lambda / 29), l = (0, 0, 1), u_i ~ U[0, 1) from MT19937 seed 20240601
    in point order (numpy RandomState.random_sample); p, n, area, rayEpsilon from the tessellator's surface points."""
    n = len(points)
    out = np.zeros(n, dt_irr)
    out["p"] = points["p"]; out["n"] = points["n"]; out["area"] = points["area"]; out["rayEpsilon"] = points["rayEpsilon"]
    u = np.random.RandomState(20240601).random_sample(n).astype(np.float32)
    base = np.maximum(np.float32(0), points["n"][:, 2]) * (np.float32(0.5) + u)
    lam = (np.float32(1) + np.float32(0.1) * np.arange(30, dtype=np.float32) / np.float32(29))
    out["E"] = base[:, None] * lam[None, :]
    return out


def random_hit_points(points, n, seed=7):
    """The incoherent hit set: n surface points drawn uniformly (with replacement) from the point set, MT19937 seed 7."""
    return np.ascontiguousarray(points["p"][np.random.RandomState(seed).randint(0, len(points), n)])


def aggregate_test_rays(bmin, bmax, n, seed=5489):
    """The ray distribution of the reference's AggregateTest (renderers/aggregatetest.cpp:64-88) without its serial
    `lastHit` chain: origins uniform in the scene bound grown by its largest extent, directions uniform on the sphere with
    1/32 of them snapped to an axis, mint 0 or 1e-3.  Rows: o, d, mint, maxt."""
    r = np.random.RandomState(seed)
    bmin = np.asarray(bmin, np.float32); bmax = np.asarray(bmax, np.float32)
    ext = (bmax - bmin).max()
    lo, hi = bmin - ext, bmax + ext
    rays = np.zeros((n, 8), np.float32)
    rays[:, :3] = lo + r.random_sample((n, 3)).astype(np.float32) * (hi - lo)
    u1, u2 = r.random_sample(n), r.random_sample(n)
    z = 1.0 - 2.0 * u1; rr = np.sqrt(np.maximum(0.0, 1.0 - z * z)); phi = 2.0 * np.pi * u2
    d = np.stack([rr * np.cos(phi), rr * np.sin(phi), z], 1).astype(np.float32)
    snap = r.randint(0, 32, n)
    for k, keep in ((0, 2), (1, 1), (2, 0)):                 # dir.x = dir.y = 0 / x = z = 0 / y = z = 0
        m = snap == k
        dd = np.zeros_like(d[m]); dd[:, keep] = np.where(d[m][:, keep] >= 0, 1, -1)
        d[m] = dd
    rays[:, 3:6] = d
    rays[:, 6] = np.where(r.random_sample(n) < 0.25, np.float32(1e-3), np.float32(0))
    rays[:, 7] = np.inf
    return rays


def through_mesh_rays(P, n, seed=11):
    """Incoherent rays that all reach the mesh: origins uniform on a sphere of twice the mesh's bounding radius, each aimed at
    a randomly chosen vertex (MT19937 seed 11).  Rows: o, d, mint, maxt."""
    r = np.random.RandomState(seed)
    c = (P.min(axis=0) + P.max(axis=0)) * 0.5
    rad = np.linalg.norm(P - c, axis=1).max()
    u1, u2 = r.random_sample(n), r.random_sample(n)
    z = 1.0 - 2.0 * u1; rr = np.sqrt(np.maximum(0.0, 1.0 - z * z)); phi = 2.0 * np.pi * u2
    o = c + 2.0 * rad * np.stack([rr * np.cos(phi), rr * np.sin(phi), z], 1)
    t = P[r.randint(0, len(P), n)].astype(np.float64)
    d = t - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((n, 8), np.float32)
    rays[:, :3] = o; rays[:, 3:6] = d; rays[:, 7] = np.inf
    return rays
