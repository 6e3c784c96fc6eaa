"""
Pins of the target / target guide restatements of the oracle (oracle/components.h) against
independent float64 numpy statements of the shader math and against the identities the shapes
must satisfy (sh/target.*.glsl, sh/target_guide.*.glsl of the reference). CPU only.
"""

import numpy as np
import pytest

import harness as H  # noqa: F401  (puts the repo root on sys.path)
import oracle
from theia_b200 import target
from theia_b200.scene import Transform

RNG = np.random.default_rng(7)
N = 4096


def unit(v):
    v = np.asarray(v, dtype=np.float64)
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


def frameOf(shape):
    """(objToWorld 3x3 with columns x y z, centre) as the shader sees it"""
    p = shape._firstBlock(0)
    m = np.array(p._objToWorld.m, dtype=np.float64).reshape(3, 3).T  # GLSL column major
    return m, np.array([p.position.x, p.position.y, p.position.z], dtype=np.float64)


TARGETS = [
    pytest.param(lambda: target.FlatTarget(width=1.2, length=0.8, position=(-1.0, 0.1, 0.0), direction=(1.0, 0.2, -0.1),
                                           up=(0.0, 0.0, 1.0)), id="flat"),
    pytest.param(lambda: target.DiskTarget(position=(-1.0, 0.0, 0.1), radius=0.6, normal=(0.9, -0.3, 0.2)), id="disk"),
]


@pytest.mark.parametrize("make", TARGETS)
def test_plane_target_frame_matches_reference_view_matrix(make):
    # src/theia/target.py:296-306 / 414-424: objToWorld = Transform.View(direction, up).innerMatrix,
    # normal = its third column, prob = 1 / area
    t = make()
    t.update(0)
    m, _ = frameOf(t)
    assert np.allclose(m.T @ m, np.eye(3), atol=1e-6)
    p = t._firstBlock(0)
    n = np.array([p._normal.x, p._normal.y, p._normal.z])
    assert np.allclose(m[:, 2], n, atol=1e-7)
    if isinstance(t, target.FlatTarget):
        assert np.allclose(n, unit(t.direction), atol=1e-6)
        assert p._prob == pytest.approx(1.0 / (1.2 * 0.8), rel=1e-6)
        ref = Transform.View(direction=t.direction, up=t.up).innerMatrix
    else:
        assert np.allclose(n, unit(t.normal), atol=1e-6)
        assert p._prob == pytest.approx(1.0 / (np.pi * 0.36), rel=1e-6)
        ref = Transform.View(direction=t.normal, up=(0.0, 1.0, 0.0)).innerMatrix
    assert np.allclose(m, ref, atol=1e-6)


@pytest.mark.parametrize("make", TARGETS)
def test_plane_target_sample_and_intersect_agree(make):
    t = make()
    t.update(0)
    m, c = frameOf(t)
    obs = (RNG.normal(size=(N, 3)) * 2.0).astype(np.float32)
    s = oracle.targetSample(t, obs)
    # samples lie on the shape, in both frames
    local = (s["position"].astype(np.float64) - c) @ m
    assert np.abs(local[:, 2]).max() < 1e-5
    assert np.allclose(local, s["objPosition"], atol=1e-5)
    if isinstance(t, target.FlatTarget):
        assert np.abs(local[:, 0]).max() <= 0.6 + 1e-6 and np.abs(local[:, 1]).max() <= 0.4 + 1e-6
        # uniform over the rectangle: first two moments
        assert abs(local[:, 0].mean()) < 0.03 and local[:, 0].var() == pytest.approx(1.2**2 / 12, rel=0.08)
        assert abs(local[:, 1].mean()) < 0.03 and local[:, 1].var() == pytest.approx(0.8**2 / 12, rel=0.08)
    else:
        r = np.hypot(local[:, 0], local[:, 1])
        assert r.max() <= 0.6 + 1e-6
        assert (r**2).mean() == pytest.approx(0.6**2 / 2, rel=0.05)  # uniform over the disk
    # the normal opposes... faces the observer; object normal is +-z
    toObs = obs.astype(np.float64) - s["position"]
    assert np.all(np.einsum("ij,ij->i", toObs, s["normal"]) > 0)
    assert np.all(np.abs(s["objNormal"][:, 2]) == 1.0) and np.all(s["objNormal"][:, :2] == 0.0)
    assert np.allclose(s["dist"], np.linalg.norm(toObs, axis=1), rtol=1e-5)
    # a ray from the observer towards the sample intersects the shape at the sample
    d = unit(-toObs).astype(np.float32)
    h = oracle.targetIntersect(t, obs, d)
    inner = (np.abs(local[:, 0]) < 0.55) & (np.abs(local[:, 1]) < 0.35) if isinstance(t, target.FlatTarget) else r < 0.55
    inner &= np.abs(np.einsum("ij,ij->i", d, s["normal"])) > 0.2  # grazing rays amplify the rounding of d
    assert np.all(h["valid"][inner] == 1.0)
    assert np.allclose(h["position"][inner], s["position"][inner], atol=1e-4)
    assert np.allclose(h["dist"][inner], s["dist"][inner], rtol=1e-4)
    assert np.allclose(h["normal"][inner], s["normal"][inner], atol=1e-6)
    assert np.allclose(h["prob"][inner], s["prob"][inner])
    # rays pointing away miss (prob 0), and planes never report valid hits behind the observer
    away = oracle.targetIntersect(t, obs, -d)
    assert np.all(away["valid"] == 0.0) and np.all(away["prob"] == 0.0)


GUIDES = [
    pytest.param(lambda: target.FlatTargetGuide(width=3.0, height=2.0, position=(0.0, -1.2, 0.3), normal=(0.0, -1.0, 0.0),
                                                up=(0.0, 0.0, 1.0)), id="flat"),
    pytest.param(lambda: target.DiskTargetGuide(position=(0.0, -1.2, 0.0), radius=1.6, normal=(0.0, -1.0, 0.0)), id="disk"),
    pytest.param(lambda: target.SphereTargetGuide(position=(0.0, 0.3, 0.0), radius=1.5), id="sphere"),
]


@pytest.mark.parametrize("make", GUIDES)
def test_guide_sample_and_eval_agree(make):
    # evalTargetGuide(dir sampled by sampleTargetGuide) must return the probability and distance of the sample
    g = make()
    g.update(0)
    obs = np.tile(np.array([[0.4, -5.0, 0.2]], np.float32), (N, 1)) + RNG.normal(size=(N, 3)).astype(np.float32) * 0.5
    s = oracle.guideSample(g, obs)
    assert np.allclose(np.linalg.norm(s["dir"], axis=1), 1.0, atol=1e-5)
    assert np.all(s["prob"] > 0)
    e = oracle.guideEval(g, obs, s["dir"])
    if isinstance(g, target.SphereTargetGuide):
        inside = np.ones(N, bool)
    else:
        # samples on the very rim may fall outside after the round trip through the direction
        inside = e["prob"] > 0
        assert inside.mean() > 0.995
    assert np.allclose(e["prob"][inside], s["prob"][inside], rtol=2e-4)
    assert np.allclose(e["dist"][inside], s["dist"][inside], rtol=2e-4)


@pytest.mark.parametrize("make", GUIDES)
def test_guide_probability_is_normalised(make):
    # int p(w) dw = 1 over the sphere of directions: Monte Carlo with uniform directions, float64 statement
    g = make()
    g.update(0)
    n = 400000
    obs0 = np.array([0.4, -5.0, 0.2], np.float32)
    w = unit(RNG.normal(size=(n, 3))).astype(np.float32)
    e = oracle.guideEval(g, np.tile(obs0, (n, 1)), w)
    p = e["prob"].astype(np.float64)
    frac = np.count_nonzero(p) / n
    integral = 4.0 * np.pi * p.mean()
    err = 4.0 / np.sqrt(max(frac * n, 1.0))
    assert integral == pytest.approx(1.0, abs=max(err, 0.02))
    # seen from behind, plane guides offer nothing; the sphere guide is two sided
    behind = np.array([0.4, 4.0, 0.2], np.float32)
    eb = oracle.guideEval(g, np.tile(behind, (n // 10, 1)), w[: n // 10])
    if isinstance(g, target.SphereTargetGuide):
        assert eb["prob"].max() > 0
    else:
        assert eb["prob"].max() == 0.0
