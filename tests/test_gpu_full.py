"""Parity at the sizes that are BENCHMARKED (BASELINE.json configs[1] and configs[4]), against the real reference:

  tests/golden/skin_full.npz   scenes/skin.scene in full -- 24.6 M irradiance points, 1280x720x4 spp, desiredLength 512 --
                               rendered by oracle/_ref/ref_driver at --ncores 1 (22 CPU-minutes): the image, the octree root,
                               every 4099th irradiance point, every 997th camera sample (hit, Mo, subsurface term, full Li)
                               and the 30 profile tables (every 16th entry + 2048 entries at both ends)
  tests/golden/mpc_kat512.npz  four two-layer thin-epidermis stacks at desiredLength 512 (2048^2 transforms)

Tolerances are north_star's: profile tables 1e-5, irradiance and Mo 1e-4 per point, image RMSE 1e-3 per channel.  Li per
sample 1e-4 on every sampled row (an allow-list for silhouette flips of single direct-lighting terms exists and is empty:
the device restates libm's float functions bit for bit, measured max 1.4e-6).
"""
import os

import numpy as np
import pytest

import golden_util as gu

ROOT = gu.ROOT
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def full():
    z = np.load(os.path.join(ROOT, "tests", "golden", "skin_full.npz"))
    meta = dict(zip([str(k) for k in z["meta_keys"]], z["meta_vals"]))
    return z, meta


@pytest.fixture(scope="module")
def frame(pkg, full):
    """skin.scene through the product path: tables, tessellation, irradiance pass, octree, sampler stream -- all on the device."""
    z, meta = full
    dev = pkg.SkinDevice(0)
    hs = pkg.HostScene(os.path.join(ROOT, "scenes", "skin.scene"))
    hs.upload(dev)
    hdr, rd = hs.build_profile(dev)
    n_pts = dev.tessellate_resident(hs.minsampledistance, False, material_id=0)
    dev.irradiance_resident(0, n_pts, hs.irradiance_tasks(n_pts, 1), hs.mix)
    n_nodes, root, bounds = dev.build_octree()
    n = dev.frame_setup(hs.frame_desc(n_cores=1))
    dev.frame_samples()
    yield dict(dev=dev, hs=hs, n=n, n_pts=n_pts, hdr=hdr, rd=rd, root=root, bounds=bounds)
    hs.close()
    dev.close()


def _table_err(got, ref):
    """max over entries of |got - ref| / |ref| with the floor SURVEY.md 8d C5 states (entries below 1e-30 of the table's maximum)."""
    ref = ref.astype(np.float64); got = got.astype(np.float64)
    floor = np.abs(ref).max(axis=-1, keepdims=True) * 1e-30
    return float((np.abs(got - ref) / np.maximum(np.abs(ref), floor)).max())


def test_profile_tables_512_vs_reference(frame, full):
    """The 30 tables skc_profile_batch builds for skin.scene (119 766 entries each) against MPC_ComputeDiffusionProfile +
    MPC_ResampleForUniformDistanceSquaredDistribution of the reference (core/multipole.cpp:241-295)."""
    z, meta = full
    rd, hdr = frame["rd"], frame["hdr"]
    assert rd.shape == (30, int(meta["profile_len"]))
    st, edge = int(z["profile_stride"]), z["profile_rd_head"].shape[1]
    errs = [_table_err(rd[:, ::st], z["profile_rd_strided"]), _table_err(rd[:, :edge], z["profile_rd_head"]), _table_err(rd[:, -edge:], z["profile_rd_tail"])]
    print("profile tables at desiredLength 512: max rel (strided, head, tail) =", errs)
    assert max(errs) < 1e-5
    assert np.allclose(hdr[:, 1], z["profile_hdr"][:, 1], rtol=1e-6, atol=0)            # rcpDsqSpacing
    assert np.allclose(hdr[:, 2:], z["profile_hdr"][:, 2:], rtol=1e-5, atol=1e-12)       # total reflectance / transmittance


def test_profile_batch_512_known_answers(pkg, frame):
    """BASELINE.json configs[4] at the size that is timed: thin two-layer stacks at desiredLength 512, d2 / R / T and totals."""
    dev = frame["dev"]
    z = np.load(os.path.join(ROOT, "tests", "golden", "mpc_kat512.npz"))
    i = 0
    while "case%d_spec" % i in z.files:
        s = z["case%d_spec" % i]
        nl, dl, lerp, res = int(s[0]), int(s[1]), int(s[2]), int(s[3])
        layers = s[5:].reshape(nl, 4).astype(np.float32)
        tot = np.float32(0)
        for l in layers:
            tot = np.float32(tot + np.float32(1) / np.float32(l[2] + l[3]))
        step = np.float32(np.float32(np.float32(12) * np.float32(tot / np.float32(nl))) / np.float32(dl))     # core/multipole.cpp:262-268
        specs = np.array([(l[0], l[1], l[2], l[3], 0.0) for l in layers], pkg.skc.DT_LAYER)
        opts = np.array([(step, dl, lerp)], pkg.skc.DT_MPC_OPTIONS)
        d2, R, T, totals = dev.profile_batch(specs, [nl], opts, resample=bool(res))
        body = np.stack([d2[0], R[0], T[0]])
        assert body.shape[1] == int(z["case%d_len" % i])
        st, edge = 16, z["case%d_d2_R_T_head" % i].shape[1]
        for name, got in (("strided", body[:, ::st]), ("head", body[:, :edge]), ("tail", body[:, -edge:])):
            ref = z["case%d_d2_R_T_%s" % (i, name)]
            assert np.array_equal(got[0], ref[0]), "d2 grid differs (%s)" % name
            e = max(_table_err(got[1], ref[1]), _table_err(got[2], ref[2]))
            assert e < 1e-5, (i, name, e)
        assert np.allclose(totals[0], z["case%d_totals" % i], rtol=1e-5)
        i += 1
    assert i == 4


def test_fft_rows_2048_equals_generic_kernel(pkg, frame, monkeypatch):
    """k_fft_rows2048 (two shared-memory crossings per row) does the butterflies of k_fft_rows in the same order: identical tables."""
    dev = frame["dev"]
    z = np.load(os.path.join(ROOT, "tests", "golden", "mpc_kat512.npz"))
    s = z["case1_spec"]
    nl, dl, lerp = int(s[0]), int(s[1]), int(s[2])
    layers = s[5:].reshape(nl, 4).astype(np.float32)
    specs = np.array([(l[0], l[1], l[2], l[3], 0.0) for l in layers], pkg.skc.DT_LAYER)
    opts = np.array([(np.float32(0.01), dl, lerp)], pkg.skc.DT_MPC_OPTIONS)
    out = {}
    for mode in ("generic", "tiled"):
        if mode == "generic":
            monkeypatch.setenv("SKC_FFT_ROWS", "generic")
        else:
            monkeypatch.delenv("SKC_FFT_ROWS", raising=False)
        out[mode] = dev.profile_batch(specs, [nl], opts, resample=False)
    for a, b in zip(out["generic"], out["tiled"]):
        assert np.array_equal(np.asarray(a).view(np.uint32), np.asarray(b).view(np.uint32))


def test_irradiance_points_vs_reference(frame, full):
    z, meta = full
    dev = frame["dev"]
    assert frame["n_pts"] == int(meta["irradiance_points"])
    ref = z["irradiance_points"]                      # p[3] n[3] E[30] area rayEpsilon
    st = int(z["irradiance_stride"])
    pa, E = dev.irradiance_rows(0, len(ref), st)
    assert np.array_equal(pa[:, :3].view(np.uint32), ref[:, :3].view(np.uint32)), "point positions differ"
    assert np.array_equal(pa[:, 3].view(np.uint32), ref[:, 36].view(np.uint32)), "point areas differ"
    e = gu.row_rel_err(E, ref[:, 6:36])
    lit = ref[:, 6:36].max(axis=1) > 0
    assert np.array_equal(E.max(axis=1) > 0, lit)
    print("irradiance: %d sampled points, %d lit, max rel %.3g" % (len(ref), lit.sum(), e.max()))
    assert e.max() < 1e-6, "measured: bit-exact (north_star allows 1e-4)"


def test_octree_root_vs_reference(frame, full):
    z, meta = full
    assert np.array_equal(frame["bounds"].view(np.uint32), z["octree_bounds"].view(np.uint32))
    r, ref = frame["root"], z["octree_root"]
    # root = sums over 24.6 M points in the tree's own order; the reference's leaf order is its insertion order as well
    assert np.array_equal(r.view(np.uint32), ref.view(np.uint32)), "root {p, sumArea, Et} over 24.6 M points: bit-exact"


def test_mo_vs_reference_at_reference_hits(frame, full):
    z, meta = full
    dev, hs = frame["dev"], frame["hs"]
    hits = z["hits"].view(gu.DT_HIT).reshape(-1)
    s = np.nonzero(hits["hasBssrdf"] == 1)[0]
    mo = dev.mo(hits["p"][s], hs.maxerror)
    e = gu.row_rel_err(mo, z["mo"][s])
    print("Mo at %d reference hit points of the full scene: max rel %.3g, mean %.3g; table: %s" % (len(s), e.max(), e.mean(), dev.mo_table_stats()))
    assert dev.mo_table_stats()["rows"] > 0, "the benchmarked kernel (compact table) must be the one tested"
    assert e.max() < 1e-5, "measured 1.4e-6 (north_star allows 1e-4)"


# Samples whose full Li may exceed 1e-4: visibility flips of single direct-lighting terms at silhouettes (see the module text).
# Counted: the test fails if there are more than this many among the 3 706 sampled, or if any of them is off by more than one
# direct-lighting term can explain (1/8 of the direct part: 2 lights x 4 samples, each a light term and a BSDF term).
LI_ALLOW = 0


def test_frame_full_vs_reference(frame, full):
    z, meta = full
    dev, hs, n = frame["dev"], frame["hs"], frame["n"]
    assert n == int(meta["camera_samples"])
    st = int(z["sample_stride"])
    xy, _, rays = dev.frame_get_samples(0, n)
    assert np.array_equal(xy[::st].view(np.uint32), z["camera_xy"].view(np.uint32)), "sampler stream differs"
    assert np.array_equal(rays[::st, :6].view(np.uint32), z["camera_rays"][:, :6].view(np.uint32)), "camera rays differ"
    dev.frame_film_clear()
    Lr = dev.frame_render(0, n, hs.maxerror, hs.mix, want_L=True)
    order = dev.frame_order(0, n)
    L = np.empty_like(Lr); L[order] = Lr
    L = L[::st]
    ref = z["L"]
    lit = ref.max(axis=1) > 0
    assert np.array_equal(L.max(axis=1) > 0, lit), "different sets of samples carry radiance"
    e = gu.row_rel_err(L, ref)
    bad = np.nonzero(e >= 1e-4)[0]
    print("full Li: %d sampled, %d lit, max rel %.3g, beyond 1e-4: %d (allowed %d)" % (len(ref), lit.sum(), e.max(), len(bad), LI_ALLOW))
    print("  rows beyond 1e-4:", [(int(i), float(e[i])) for i in bad])
    assert len(bad) <= LI_ALLOW
    d_ref = np.abs(ref - z["L_sss"]).max(axis=1)
    for i in bad:       # a flipped term changes the row by at most its share of the direct part
        assert np.abs(L[i] - ref[i]).max() <= 0.3 * d_ref[i] + 1e-4 * np.abs(ref[i]).max(), (i, e[i])
    # the image (film/image.cpp:178-213), RMSE per channel, and its row maxima
    rgb = hs.film_to_rgb(dev.frame_film())
    img = z["image"]
    assert rgb.shape == img.shape == (720, 1280, 3)
    for c in range(3):
        rmse = np.sqrt(np.mean((rgb[..., c] - img[..., c]) ** 2)); scale = np.sqrt(np.mean(img[..., c] ** 2))
        print("channel %d: RMSE / RMS = %.3g" % (c, rmse / scale))
        assert rmse < 1e-3 * scale
    assert np.allclose(rgb.max(axis=1), img.max(axis=1), rtol=1e-3, atol=1e-4 * img.max())
