"""The irradiance kernel does not draw light samples for an emitter sphere that lies wholly below a point's horizon
(csrc/skc_irradiance.cu k_irr_sample): its claim is that the reference would have discarded every such sample at
`if (Dot(wi, n) <= 0.) continue` (integrators/multipolesubsurface.cpp:121) whatever the sample values are.  Checked here on the
CPU against the C restatement of IrradianceTask::Run (oracle/skin_oracle.c, pinned bit-exact to the real reference by
tests/test_oracle.py): with every other light switched off, each point for which the kernel's predicate -- evaluated in numpy
in the kernel's fp32 operation order -- holds must come out of the reference's own loop with E == 0 in all 30 samples.  The
points are the head's own surface points and a cloud of synthetic points placed around the margin of the predicate."""
import numpy as np
import pytest

import golden_util as gu
import oracle_api as orc

F = np.float32


def _below_horizon(p, n, light):
    """The kernel's test, fp32, same operation order: dot3(toCentre, n) <= -1.01f * radius - 1e-5f * reach."""
    c = light[[3, 7, 11]].astype(F)                     # ObjectToWorld(Point(0, 0, 0)): the translation column of o2w
    radius = F(light[32])
    t = (c[None, :] - p).astype(F)
    reach = (np.abs(t[:, 0]) + np.abs(t[:, 1])).astype(F) + np.abs(t[:, 2])
    d = ((t[:, 0] * n[:, 0]).astype(F) + (t[:, 1] * n[:, 1]).astype(F)).astype(F) + (t[:, 2] * n[:, 2]).astype(F)
    return d.astype(F) <= (F(-1.01) * radius - (F(1e-5) * reach).astype(F)).astype(F)


@pytest.fixture(scope="module")
def tiny():
    return gu.load_scene_golden("skin_tiny")


def _only_light(g, k):
    lights = np.array(g.lights, np.float32, copy=True)
    for j in range(len(lights)):
        if j != k:
            lights[j, 40:70] = 0.0                      # Lemit = 0: Sample_L returns black, the reference skips the sample (:122)
    return lights


def test_points_of_the_head(tiny):
    g = tiny
    pts = np.array(g.surface_points, copy=True)
    p = np.ascontiguousarray(pts["p"], F); n = np.ascontiguousarray(pts["n"], F)
    for k in range(len(g.lights)):
        skip = _below_horizon(p, n, np.asarray(g.lights[k], F))
        assert 0.25 < skip.mean() < 0.75, skip.mean()   # a closed head: roughly half of it faces away from either emitter
        g2 = type(g)(**vars(g)); g2.lights = _only_light(g, k)
        out = orc.irradiance(g2)
        E = out["E"]
        assert (E[skip] == 0).all(), "light %d: %d skipped points with E != 0" % (k, int((E[skip] != 0).any(axis=1).sum()))
        assert (E[~skip] != 0).any()                    # and the test is not vacuous: the others do receive light


def test_points_around_the_margin(tiny):
    """Points in free space around emitter 0, normals tilted so that n . (c - p) sits within a few per cent of the predicate's
    threshold on both sides; distances from 1.5 to 40 radii.  Wherever the predicate says skip, the reference's loop gives 0."""
    g = tiny
    light = np.asarray(g.lights[0], F)
    c = light[[3, 7, 11]].astype(np.float64); r = float(light[32])
    rng = np.random.default_rng(11)
    m = 20000
    direction = rng.normal(size=(m, 3)); direction /= np.linalg.norm(direction, axis=1, keepdims=True)
    dist = r * rng.uniform(1.5, 40.0, m)
    p = c[None, :] - direction * dist[:, None]                               # c - p = direction * dist
    # a normal with n . direction = -(1.01 r + 1e-5 reach) / dist * f, f in [0.9, 1.1]
    reach = np.abs(direction * dist[:, None]).sum(axis=1)
    cosang = -(1.01 * r + 1e-5 * reach) / dist * rng.uniform(0.9, 1.1, m)
    perp = np.cross(direction, rng.normal(size=(m, 3))); perp /= np.linalg.norm(perp, axis=1, keepdims=True)
    n = cosang[:, None] * direction + np.sqrt(np.maximum(0.0, 1 - cosang ** 2))[:, None] * perp
    pts = np.zeros(m, gu.DT_SP)
    pts["p"] = p.astype(F); pts["n"] = n.astype(F); pts["area"] = 1.0; pts["rayEpsilon"] = 1e-4
    skip = _below_horizon(np.ascontiguousarray(pts["p"]), np.ascontiguousarray(pts["n"]), light)
    assert 0.3 < skip.mean() < 0.7, skip.mean()
    g2 = type(g)(**vars(g)); g2.lights = _only_light(g, 0)
    out = orc.irradiance(g2, points=pts, n_tasks=64)
    E = out["E"]
    assert (E[skip] == 0).all(), "%d skipped points with E != 0" % int((E[skip] != 0).any(axis=1).sum())
    lit = (E[~skip] != 0).any(axis=1)
    print("margin cloud: %d of %d points skipped; %d of the %d others see a sliver of the emitter" % (skip.sum(), m, lit.sum(), (~skip).sum()))
    assert lit.any()                                    # just above the threshold the emitter's rim does clear the horizon
