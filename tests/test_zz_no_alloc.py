"""skc_mo_static: the no-allocation variant of skc_mo (scratch in a file-scope static __device__ array; no counterpart
in the reference).  Named test_zz_* so that pytest collects it last:
a failure here must not hide the parity tests.

Checks, through the C ABI: the output equals skc_mo's on two calls in a row (the array keeps the last call's contents,
so the second call runs on dirty scratch, with a different size), and sizes beyond the stated maximum are rejected."""
import numpy as np
import pytest

import golden_util as gu

pytestmark = pytest.mark.gpu


def _octree_device(pkg, g, tables="golden64"):
    """tables: "golden64" = the golden's desiredLength-64 tables (exact-table kernel k_mo_packet); "device512" = the
    desiredLength-512 tables of scenes/skin.scene built on the device (compact shared-memory table, k_mo_lambda)."""
    import os
    dev = pkg.SkinDevice(0)
    gu.setup_device(dev, g)
    if tables == "device512":
        hs = pkg.HostScene(os.path.join(gu.ROOT, "scenes", "skin.scene"))
        try:
            hs.build_profile(dev)
        finally:
            hs.close()
        assert dev.mo_table_stats()["rows"] > 0
    dev.set_irradiance_points(g.irradiance_points)
    dev.build_octree()
    return dev


@pytest.mark.parametrize("tables", ["golden64", "device512"])
def test_mo_static_matches_mo_twice_in_a_row(pkg, tiny, tables):
    dev = _octree_device(pkg, tiny, tables)
    try:
        hp = tiny.hits["p"][tiny.hits["hasBssrdf"] == 1]
        rng = np.random.default_rng(11)          # the golden has a few hundred hits: add jittered copies (Mo takes any point)
        p = np.concatenate([hp] + [hp + rng.normal(0, 0.01, hp.shape).astype(np.float32) for _ in range(5)]).astype(np.float32)
        assert len(p) > 3000
        a, b = p[:3000], p[1000:2200][::-1]             # two different calls, the second smaller and in another order
        for pts in (a, b):
            ref = dev.mo(pts, tiny.max_error)           # existing entry point
            got = dev.mo_static(pts, tiny.max_error)    # no-allocation entry point
            assert np.isfinite(got).all() and np.abs(ref).max() > 0
            e = gu.row_rel_err(got, ref)
            assert e.max() < 1e-6, e.max()               # same kernels, same per-point summation order: expected identical
        # and against the reference's own Mo values (tolerance 1e-4 of north_star, as in test_gpu_parity)
        if tables != "golden64":
            return
        s = np.nonzero(tiny.hits["hasBssrdf"] == 1)[0]
        e = gu.row_rel_err(dev.mo_static(tiny.hits["p"][s], tiny.max_error), tiny.mo[s])
        assert e.max() < 1e-4, e.max()
    finally:
        dev.close()


def test_mo_static_rejects_sizes_beyond_the_maximum(pkg, tiny):
    dev = _octree_device(pkg, tiny)
    try:
        n = dev.MO_STATIC_MAX_HITS + 1
        with pytest.raises(pkg.SkcError, match="exceed the static scratch capacity"):
            dev.mo_static(np.zeros((n, 3), np.float32), tiny.max_error)
        full = np.resize(tiny.hits["p"][tiny.hits["hasBssrdf"] == 1], (dev.MO_STATIC_MAX_HITS, 3))
        e = gu.row_rel_err(dev.mo_static(full, tiny.max_error), dev.mo(full, tiny.max_error))     # exactly the maximum is accepted
        assert e.max() < 1e-6
    finally:
        dev.close()
