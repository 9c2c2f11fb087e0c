"""The compact shared-memory form of the Rd tables (csrc/skc_mo_table.h) against sampleProfile
(core/multipole.cpp:60-73), over ALL (wavelength, d2): every lookup the Mo kernel makes is either the
reference's exact one or within 1e-5 of it.  CPU only: the builder and evaluator are host code of libskincuda.so;
the 30 desiredLength-512 tables of scenes/skin.scene come from the C restatement of the profile calculator
(oracle/mpc_oracle.c), which tests/test_oracle.py pins to the real reference."""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import oracle_api as orc
from golden_util import ROOT

TOL = 1e-5


@pytest.fixture(scope="module")
def skin_tables(pkg):
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin.scene"))
    try:
        c = hs.skin_coefficients()
    finally:
        hs.close()

    def one(l):
        specs = [(c["eta"][k], c["thickness"][k], c["mua"][k][l], c["musp"][k][l], 0.0) for k in range(2)]
        mfp = np.float32(np.float32(np.float32(1) / np.float32(c["mua"][0][l] + c["musp"][0][l]) +
                                    np.float32(1) / np.float32(c["mua"][1][l] + c["musp"][1][l])) / np.float32(2))
        step = np.float32(np.float32(np.float32(12) * mfp) / np.float32(512))     # core/multipole.cpp:262-268
        d2, R, T, tot = orc.mpc_profile(specs, (step, 512, 1), resample=True)
        return R, np.float32(np.float32(len(d2) - 1) / d2[-1])                    # rcpDsqSpacing, multipole.cpp:274-275

    with ThreadPoolExecutor(8) as ex:          # ctypes releases the GIL
        res = list(ex.map(one, range(30)))
    table = np.stack([r[0] for r in res]).astype(np.float32)
    rcp = np.array([r[1] for r in res], np.float32)
    assert table.shape == (30, 119766)
    return table, rcp


def test_compact_table_bound_over_all_lookups(pkg, skin_tables):
    table, rcp = skin_tables
    cp = pkg.skc.CompactProfile(table, rcp, tol=TOL)
    assert cp.h, "no wavelength got a compact window"
    try:
        L = table.shape[1]
        assert cp.rows * 128 + cp.head_floats * 4 <= 110 * 1024          # shared-memory budget beside 16 warps' traversal state
        assert cp.coverage > 0.75, cp.coverage                            # log-weighted share of lookups served from shared memory
        assert (cp.f_hi <= L - 2).all() and (cp.head_len <= cp.f_lo).all()
        rng = np.random.default_rng(5)
        worst = 0.0
        for l in range(30):
            i = np.arange(L + 40, dtype=np.float64)                        # past the table's end too: exactly 0 there
            f = np.concatenate([i + o for o in (0.0, 0.13, 0.27, 0.5, 0.77, 0.93)] + [rng.uniform(0, L - 1, 200000)])
            d2 = (f / np.float64(rcp[l])).astype(np.float32)
            out, ex = cp.eval(l, d2)
            fd = d2 * rcp[l]                                               # the kernel's own f (fp32 product)
            inwin = (fd >= cp.f_lo[l]) & (fd < cp.f_hi[l])
            assert np.array_equal(out[~inwin], ex[~inwin])                 # outside the window: the exact rule, bit for bit
            assert (out[fd >= np.float32(L - 1)] == 0).all()               # "ensure integral convergence" cut-off
            e = np.abs(out[inwin].astype(np.float64) - ex[inwin]) / np.abs(ex[inwin].astype(np.float64))
            worst = max(worst, float(e.max()))
        assert worst <= TOL, worst
    finally:
        cp.close()


def test_short_tables_stay_exact(pkg):
    # desiredLength 64 tables (2 324 entries) are too coarse for a window worth having: the builder says so and
    # skc_upload_profile keeps the exact kernel
    z = np.load(os.path.join(ROOT, "tests", "golden", "skin_tiny.npz"))
    cp = pkg.skc.CompactProfile(z["profile_rd"], z["profile_hdr"][:, 1], tol=TOL)
    assert not cp.h
