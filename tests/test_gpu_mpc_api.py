"""B4: the reference's MPC_* DLL API (include/mpc_cuda.h, libMultipoleProfileCalculator.so) driven the way its call sites do
(host/mpc/mpc_driver.cpp mirrors core/multipole.cpp:241-295 and renderers/profilefit.cpp:190-196), against the real
reference's outputs for the same calls (tests/golden/mpc_api.npz, mpc_kat.npz: ref_driver mpc)."""
import os
import subprocess

import numpy as np
import pytest

import golden_util as gu

PKG = os.path.join(gu.ROOT, "pbrt-v2-skin_b200")


def _cases(npz):
    z = np.load(os.path.join(gu.GOLDEN, npz))
    out, i = [], 0
    while "case%d_spec" % i in z.files:
        out.append((z["case%d_spec" % i], z["case%d_totals" % i], z["case%d_d2_R_T" % i]))
        i += 1
    return out


def test_library_exports_the_reference_api():
    """CPU: the shim library exists and exports the six MPC_* functions with the reference's C++ signatures."""
    lib = os.path.join(PKG, "libMultipoleProfileCalculator.so")
    assert os.path.exists(lib), "make -C pbrt-v2-skin_b200"
    syms = subprocess.run(["nm", "-DC", lib], stdout=subprocess.PIPE, text=True, check=True).stdout
    for sig in ("MPC_CreateOutput(unsigned int, MPC_Output**)",
                "MPC_ComputeDiffusionProfile(unsigned int, MPC_LayerSpec const*, MPC_Options const*, MPC_Output**)",
                "MPC_ResampleForUniformDistanceSquaredDistribution(MPC_Output*, unsigned int)",
                "MPC_ResampleDistribution(MPC_Output*, unsigned int, float*)", "MPC_FreeOutput(MPC_Output*)", "MPC_ClearCache()"):
        assert " T " + sig in syms, sig


@pytest.mark.gpu
@pytest.mark.parametrize("npz", ["mpc_api.npz", "mpc_kat.npz"])
def test_mpc_api_matches_the_reference(npz, tmp_path):
    cases = _cases(npz)
    inp, outp = tmp_path / "in.txt", tmp_path / "out.bin"
    with open(inp, "w") as f:
        for spec, _, _ in cases:
            nl = int(spec[0])
            f.write("%d %d %d %d %r " % (nl, int(spec[1]), int(spec[2]), int(spec[3]), float(spec[4])) + " ".join(repr(float(x)) for x in spec[5:5 + 4 * nl]) + "\n")
    r = subprocess.run([os.path.join(PKG, "mpc_driver"), str(inp), str(outp)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    assert r.stdout.strip() == "mpc_profiles %d" % len(cases)
    blob, off = open(outp, "rb").read(), 0
    for k, (spec, tot, ref) in enumerate(cases):
        ln = int(np.frombuffer(blob, np.uint32, 1, off)[0]); off += 4
        got_tot = np.frombuffer(blob, np.float32, 2, off); off += 8
        got = np.frombuffer(blob, np.float32, 3 * ln, off).reshape(3, ln); off += 12 * ln
        assert ln == ref.shape[1], (k, ln, ref.shape)
        assert np.array_equal(got[0].view(np.uint32), ref[0].view(np.uint32)), "case %d: distances differ" % k
        for row in (1, 2):
            scale = np.abs(ref[row]).max()
            err = np.abs(got[row].astype(np.float64) - ref[row]) / np.maximum(np.abs(ref[row]), 1e-30 * max(scale, 1e-30))
            assert err.max() < 1e-5, (k, row, err.max())          # north_star: profile tables within 1e-5 relative
        assert np.allclose(got_tot, tot, rtol=1e-5)
    assert off == len(blob)
