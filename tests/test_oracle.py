"""Pins the C restatement (oracle/skin_oracle.c) against the REAL reference's exports
(tests/golden/skin_tiny.npz from oracle/_ref/ref_driver): the oracle may only be trusted as a
checker for the CUDA path because these pass.  CPU only."""
import numpy as np

import golden_util as gu
import oracle_api as orc


def test_closest_hit_bit_exact(tiny):
    hits = orc.trace_closest(tiny, tiny.camera_rays)
    ref = tiny.hits
    assert np.array_equal(hits["hit"], ref["hit"]) and np.array_equal(hits["prim"], np.where(ref["hit"] == 1, ref["prim"], -1))
    m = ref["hit"] == 1
    for f in ("t", "rayEpsilon", "wo"):
        assert np.array_equal(hits[f][m].view(np.uint32), ref[f][m].view(np.uint32)), f
    s = ref["hasBssrdf"] == 1
    assert np.array_equal(hits["hasBssrdf"], ref["hasBssrdf"])
    assert np.array_equal(hits["p"][s].view(np.uint32), ref["p"][s].view(np.uint32))
    assert np.array_equal(hits["n"][s].view(np.uint32), ref["n"][s].view(np.uint32))
    assert np.array_equal(orc.trace_any(tiny, tiny.camera_rays).astype(np.int32), ref["hit"])


def test_irradiance_bit_exact(tiny):
    out = orc.irradiance(tiny)
    ref = tiny.irradiance_points
    assert np.array_equal(out["p"].view(np.uint32), ref["p"].view(np.uint32))
    exact = (out["E"].view(np.uint32) == ref["E"].view(np.uint32)).all(1)
    e = gu.row_rel_err(out["E"], ref["E"])
    print("oracle irradiance: bit-exact rows %.5f, max rel %.3g" % (exact.mean(), e.max()))
    assert exact.mean() > 0.999 and e.max() < 1e-5


def test_octree_and_mo(tiny):
    t = orc.Octree(tiny.irradiance_points, tiny.cie[1])
    n, root, bounds = t.info()
    assert np.array_equal(bounds.view(np.uint32), tiny.octree_bounds.view(np.uint32))
    assert np.array_equal(root.view(np.uint32), tiny.octree_root.view(np.uint32))
    s = tiny.hits["hasBssrdf"] == 1
    mo = t.mo(tiny.hits["p"][s], tiny.profile_rd, tiny.profile_hdr[:, 1], tiny.max_error)
    assert np.array_equal(mo.view(np.uint32), tiny.mo[s].view(np.uint32)), "Mo not bit-exact"
    full = np.zeros_like(tiny.mo); full[s] = mo
    L = orc.li_sss(tiny.hits, full, tiny.rho_hd[:1025], tiny.mix)
    assert np.array_equal(L.view(np.uint32), tiny.L_sss.view(np.uint32)), "Li SSS term not bit-exact"


def test_scrambles_stream_layout():
    # 3 draws per (point, light) in point order inside each task slice, RNG(taskNum*47)
    s = orc.scrambles(1000, 32, 2)
    assert s.shape == (1000, 2, 3) and len(np.unique(s)) > 5900
    # task 0 seeds RNG(0): first MT19937 output for seed 0
    assert s[0, 0, 0] == 2357136044


def test_mpc_oracle_pinned_to_reference_known_answers():
    """oracle/mpc_oracle.c against MPC_ComputeDiffusionProfile of the real reference (tests/golden/mpc_kat.npz):
    lengths and d2 abscissae bit-exact; R, T within 1e-6 of the table's own scale (own FFT: the rounding noise of the
    transform differs, the dipole grids do not)."""
    from golden_util import load_mpc_cases
    for k, (specs, opt, res, totals, ref) in enumerate(load_mpc_cases()):
        d2, R, T, tot = orc.mpc_profile(specs, opt, resample=bool(res))
        assert len(d2) == ref.shape[1], "case %d" % k
        assert np.array_equal(d2, ref[0]), "case %d d2" % k
        for got, want, name in ((R, ref[1], "R"), (T, ref[2], "T")):
            floor = np.abs(want).max() * 1e-9
            e = np.abs(got.astype(np.float64) - want) / np.maximum(np.abs(want), floor)
            assert e.max() < 1e-6, "case %d %s: %.3g" % (k, name, e.max())
        assert np.allclose(tot, totals, rtol=1e-6, atol=1e-12)
