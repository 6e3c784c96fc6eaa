"""
Ports of the reference's shader unit tests: every numpy pin the reference's own test suite holds for
the functions on the forward path, with the reference's inputs, seeds and tolerances, evaluated through
the unit seam (theia_unit_eval / oracle/units.h) and the sampler stages (LightSampler, TargetSampler,
TargetGuideSampler).

Each test runs twice: backend "oracle" = the CPU restatement (pins the oracle to the reference's
expectations; runs without a GPU), backend "cuda" = the device functions the tracing kernels use
(`-m gpu`). Reference files, all under /root/reference/tests: test_scatter.py, test_polarization.py,
test_ray.py, test_math.py, test_random.py, test_target.py, test_target_guide.py, test_light.py.
"""

import numpy as np
import pytest
from scipy.spatial.transform import Rotation as R

import harness as H  # noqa: F401  (puts the repo root on sys.path)
from theia_b200 import _abi, _device
import theia_b200.units as u

BACKENDS = [pytest.param("oracle", id="oracle"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]
pytestmark = pytest.mark.parametrize("backend", BACKENDS)


def device(backend):
    return _device.useDevice("cpu" if backend == "oracle" else "cuda")


def ev(backend, unit, rows, n=None, **kw):
    if backend == "oracle":
        import oracle
        return oracle.evalUnit(unit, rows, n, **kw)
    from theia_b200 import _units
    return _units.evalUnit(unit, rows, n, **kw)


def runSampler(backend, sampler, i=0):
    if backend == "oracle":
        import oracle
        oracle.runSampler(sampler, i)
    else:
        import torch
        from theia_b200.pipeline import runPipeline
        runPipeline(sampler.collectStages(), i)
        torch.cuda.synchronize()


@pytest.fixture
def rng():
    return np.random.default_rng(0xC0FFEE)  # tests/conftest.py:74-76 of the reference


def mediumWords(addr, n):
    """a device address as the two float-typed words of an input row"""
    w = np.array([addr & 0xFFFFFFFF, addr >> 32], dtype=np.uint32).view(np.float32)
    return np.tile(w, (n, 1))


def unitVectors(rng, n):
    phi = rng.random(n) * 2.0 * np.pi
    cos_theta = rng.random(n) * 2.0 - 1.0
    sin_theta = np.sqrt(1.0 - cos_theta**2)
    return np.stack([sin_theta * np.cos(phi), sin_theta * np.sin(phi), cos_theta], -1)


# ------------------------------------------------------------------ tests/test_scatter.py
def test_scatterDir(backend, rng):  # tests/test_scatter.py:15-86
    N = 32 * 1024
    theta = rng.random(N) * np.pi
    phi = rng.random(N) * 2.0 * np.pi
    inDir = np.stack([np.sin(theta) * np.cos(phi), np.sin(theta) * np.sin(phi), np.cos(theta)], -1)
    cos_theta = 2.0 * rng.random(N) - 1.0
    phi_q = 2.0 * np.pi * rng.random(N)
    rows = np.column_stack([inDir, cos_theta, phi_q]).astype(np.float32)
    results = ev(backend, _abi.UNIT_SCATTER_DIR, rows).astype(np.float64)

    dots = np.multiply(rows[:, :3].astype(np.float64), results).sum(-1)
    assert np.abs(dots - rows[:, 3]).max() < 5e-7
    # phi: uniform around the incoming direction
    m = np.stack([
        np.stack([np.cos(theta) * np.cos(phi), -np.sin(phi), np.sin(theta) * np.cos(phi)], -1),
        np.stack([np.cos(theta) * np.sin(phi), np.cos(phi), np.sin(theta) * np.sin(phi)], -1),
        np.stack([-np.sin(theta), np.zeros_like(theta), np.cos(theta)], -1),
    ], -1)
    proj = np.einsum("ikj,ij->ik", m, results)
    phi_out = np.arctan2(proj[:, 1], proj[:, 0])
    bins = 32
    counts, _ = np.histogram(phi_out, bins=bins)
    assert np.abs(counts / N - 1.0 / bins).max() < 1.0 / np.sqrt(N / bins)


def reflectance(cos_i, n_i, n_t):
    """Fresnel equations (tests/test_scatter.py:89-97)"""
    sin_i = np.sqrt(np.maximum(1.0 - cos_i * cos_i, 0.0))
    sin_t = sin_i * n_i / n_t
    cos_t = np.sqrt(np.maximum(1.0 - sin_t * sin_t, 0.0))
    rs = (n_i * cos_i - n_t * cos_t) / (n_i * cos_i + n_t * cos_t)
    rp = (n_t * cos_i - n_i * cos_t) / (n_t * cos_i + n_i * cos_t)
    return rs, rp


def test_reflectance(backend, rng):  # tests/test_scatter.py:100-173
    from theia_b200 import material
    N = 32 * 1024

    class WaterModel(material.WaterBaseModel, material.MediumModel):
        def __init__(self) -> None:
            material.WaterBaseModel.__init__(self, 5.0, 1000.0, 35.0)

    with device(backend):
        inside, outside = material.BK7Model(), WaterModel()
        mat = material.Material("material", inside.createMedium(name="glass"), outside.createMedium(name="water"))
        store = material.MaterialStore([mat])
    dirs, nrms = unitVectors(rng, N), unitVectors(rng, N)
    lam = 300.0 * rng.random(N) + 400.0
    rows = np.column_stack([dirs, nrms, lam]).astype(np.float32)
    results = ev(backend, _abi.UNIT_REFLECTANCE, rows, material=store.material["material"])

    q = rows.astype(np.float64)
    cos_theta = np.multiply(q[:, :3], q[:, 3:6]).sum(-1)
    n_inside = inside.refractive_index(q[:, 6])
    n_outside = outside.refractive_index(q[:, 6])
    mask = cos_theta <= 0.0
    n_i = np.where(mask, n_outside, n_inside)
    n_t = np.where(mask, n_inside, n_outside)
    rs, rp = reflectance(np.abs(cos_theta), n_i, n_t)
    assert np.abs(rs - results[:, 0]).max() < 1e-4
    assert np.abs(rp - results[:, 1]).max() < 1e-4
    assert np.abs(n_t - results[:, 2]).max() < 1e-6


def _waterAndEmpty(backend):
    from theia_b200 import material
    from theia_b200.testing import WaterTestModel
    with device(backend):
        model = WaterTestModel()
        water = model.createMedium()
        empty = material.MediumModel().createMedium(name="empty")
        store = material.MaterialStore([], media=[water, empty])
    return model, store


def test_volumeScatter(backend, rng):  # tests/test_scatter.py:176-266
    from theia_b200 import random
    N, N_EMPTY = 32 * 1024, 32 * 64
    model, store = _waterAndEmpty(backend)
    with device(backend):
        philox = random.PhiloxRNG(key=0xC0110FFC0FFEE)
        philox.update(0)
    dir_in = unitVectors(rng, N)
    med = mediumWords(store.media["water"], N)
    med[:N_EMPTY] = mediumWords(store.media["empty"], N_EMPTY)
    rows = np.column_stack([dir_in.astype(np.float32), med])
    result = ev(backend, _abi.UNIT_VOLUME_SCATTER, rows, rng=philox)
    dir_in = rows[:, :3].astype(np.float64)
    dir_out, prob = result[:, :3].astype(np.float64), result[:, 3]

    cos_theta = np.multiply(dir_in, dir_out).sum(-1)
    bins = np.histogram(cos_theta[:N_EMPTY], bins=10, range=(-1, 1))[0]
    assert np.abs(bins / N_EMPTY - 0.1).max() < 0.1
    assert np.allclose(prob[:N_EMPTY], 1.0 / (4.0 * np.pi))
    assert np.abs(cos_theta[N_EMPTY:].mean() - model.g) < 5e-3  # Henyey-Greenstein: E[cos theta] = g
    p_exp = np.exp(model.log_phase_function(cos_theta[N_EMPTY:]))
    assert np.abs(p_exp - prob[N_EMPTY:]).max() < 5e-5
    # azimuth uniform in a local frame around the incoming direction (PBRT v4, 3.3.3)
    s = np.copysign(1.0, dir_in[:, 2])
    a = -1.0 / (s + dir_in[:, 2])
    b = a * dir_in[:, 0] * dir_in[:, 1]
    x_ax = np.stack([1.0 + s * dir_in[:, 0] ** 2 * a, s * b, -s * dir_in[:, 0]], -1)
    y_ax = np.stack([b, s + dir_in[:, 1] ** 2 * a, -dir_in[:, 1]], -1)
    x_ax /= np.sqrt(np.square(x_ax).sum(-1))[:, None]
    y_ax /= np.sqrt(np.square(y_ax).sum(-1))[:, None]
    phi = np.arctan2(np.multiply(y_ax, dir_out).sum(-1), np.multiply(x_ax, dir_out).sum(-1))
    hist = np.histogram(phi, bins=10, range=(-np.pi, np.pi))[0]
    assert np.abs((hist / N) - 0.1).max() < 0.01


def test_volumeScatterProb(backend, rng):  # tests/test_scatter.py:269-331
    N = 32 * 256
    model, store = _waterAndEmpty(backend)
    inDir, outDir = unitVectors(rng, N), unitVectors(rng, N)
    med = mediumWords(store.media["water"], N)
    med[N // 2:] = mediumWords(store.media["empty"], N - N // 2)
    rows = np.column_stack([inDir.astype(np.float32), outDir.astype(np.float32), med])
    result = ev(backend, _abi.UNIT_SCATTER_PROB, rows)[:, 0]
    cos_theta = np.multiply(rows[:, :3].astype(np.float64), rows[:, 3:6].astype(np.float64)).sum(-1)
    p = np.empty(N)
    p[: N // 2] = np.exp(model.log_phase_function(cos_theta[: N // 2]))
    p[N // 2:] = 1.0 / (4.0 * np.pi)
    assert np.abs(result - p).max() < 5e-5


# ------------------------------------------------------------------ tests/test_polarization.py
def test_polarizationRotate_phi(backend, rng):  # tests/test_polarization.py:14-61
    N = 32 * 1024
    phi = rng.random(N) * 2.0 * np.pi
    stokes = rng.random((N, 4))
    rows = np.column_stack([stokes, phi]).astype(np.float32)
    results = ev(backend, _abi.UNIT_ROTATE_PHI, rows)
    expected = stokes.copy()
    cos_phi, sin_phi = np.cos(2 * phi), np.sin(2 * phi)
    expected[:, 1] = cos_phi * stokes[:, 1] - sin_phi * stokes[:, 2]
    expected[:, 2] = sin_phi * stokes[:, 1] + cos_phi * stokes[:, 2]
    assert np.abs(results - expected).max() < 5e-6


def test_alignPolRef(backend):  # tests/test_polarization.py:64-90
    from theia_b200 import random
    N = 32 * 1024
    with device(backend):
        philox = random.PhiloxRNG(key=0xC0FFEE)
        philox.update(0)
    result = ev(backend, _abi.UNIT_ALIGN_POLREF, None, N, rng=philox).reshape((N, 4, 4))
    assert np.allclose(result, np.identity(4)[None, ...], atol=5e-6)


def test_polarizationRotate_dir(backend, rng):  # tests/test_polarization.py:93-165
    N = 32 * 1024
    phi = rng.random(N) * 2.0 * np.pi
    theta = rng.random(N) * np.pi
    stokes = rng.random((N, 4))
    dir = rng.random((N, 3))
    dir /= np.sqrt(np.square(dir).sum(-1))[:, None]
    ref = np.cross(dir, rng.random((N, 3)))
    ref /= np.sqrt(np.square(ref).sum(-1))[:, None]
    rows = np.column_stack([stokes, dir, ref, phi, theta]).astype(np.float32)
    results = ev(backend, _abi.UNIT_ROTATE_DIR, rows)
    stokes_out, ref_out = results[:, :4], results[:, 4:7]

    # scatterDir does not promise to rotate by exactly phi: it may add a constant offset per direction
    cos_phi = (ref * ref_out).sum(-1)
    sin_phi = (np.cross(ref, ref_out) * dir).sum(-1)
    phi = np.arctan2(sin_phi, cos_phi)
    c2 = 2.0 * cos_phi * cos_phi - 1.0
    s2 = 2.0 * cos_phi * sin_phi
    stokes_exp = stokes.copy()
    stokes_exp[:, 1] = c2 * stokes[:, 1] - s2 * stokes[:, 2]
    stokes_exp[:, 2] = s2 * stokes[:, 1] + c2 * stokes[:, 2]
    ref_exp = R.from_rotvec(dir * phi[:, None]).apply(ref)
    assert np.abs(ref_out - ref_exp).max() < 1e-6
    assert np.abs(stokes_out - stokes_exp).max() < 1e-6
    assert np.abs(np.square(ref_out).sum(-1).max() - 1.0) < 1e-6
    assert np.abs((ref_out * dir).sum(-1)).max() < 1e-6


def test_phaseMatrix(backend, rng):  # tests/test_polarization.py:168-255
    from theia_b200 import material
    from theia_b200.testing import WaterTestModel
    N = 32 * 1024
    with device(backend):
        water_model = WaterTestModel()  # the same model the reference test builds inline
        store = material.MaterialStore([], media=[water_model.createMedium()])
    cos_theta = rng.random(N) * 2.0 - 1.0
    stokes = rng.random((N, 4))
    rows = np.column_stack([stokes, cos_theta]).astype(np.float32)
    results = ev(backend, _abi.UNIT_PHASE_MATRIX, rows, medium=store.media["water"])

    ones, zero = np.ones(N), np.zeros(N)
    m12 = zero if (m := water_model.phase_m12(cos_theta)) is None else m
    m22 = zero if (m := water_model.phase_m22(cos_theta)) is None else m
    m33 = zero if (m := water_model.phase_m33(cos_theta)) is None else m
    m34 = zero if (m := water_model.phase_m34(cos_theta)) is None else m
    phase = np.stack([ones, m12, zero, zero, m12, m22, zero, zero, zero, zero, m33, m34, zero, zero, -m34, m33],
                     axis=-1).reshape((N, 4, 4))
    expected = np.einsum("ijk,ik->ij", phase, stokes)
    assert np.abs(results - expected).max() < 1e-3


# ------------------------------------------------------------------ tests/test_ray.py (forward)
@pytest.mark.parametrize("polarized", [True, False])
def test_surface(backend, polarized):  # tests/test_ray.py:32-263, forward=True
    from theia_b200 import light, material, random
    from theia_b200.testing import WaterTestModel
    N = 32 * 1024
    lam_min, lam_max = 200.0 * u.nm, 800.0 * u.nm
    normal = np.array([0.8, 0.36, 0.48])
    with device(backend):
        waterModel, glassModel = WaterTestModel(), material.BK7Model()
        water = waterModel.createMedium(lam_min, lam_max)
        glass = glassModel.createMedium(lam_min, lam_max)
        matStore = material.MaterialStore([material.Material("mat", glass, water)])
        philox = random.PhiloxRNG(key=0xC0FFEE)
        source = light.SphericalLightSource(timeRange=(0.0, 0.0))
        philox.update(0)
        source.update(0)
    r = ev(backend, _abi.UNIT_SURFACE, None, N, rng=philox, polarized=polarized, material=matStore.material["mat"],
           light=source, params=(*normal, lam_min, lam_max)).astype(np.float64)
    lam, rayDir, contrib = r[:, 0], r[:, 1:4], r[:, 4]
    refl_pos, refl_dir, refl_contrib = r[:, 5:8], r[:, 8:11], r[:, 11]
    trans_pos, trans_dir, trans_contrib = r[:, 12:15], r[:, 15:18], r[:, 18]

    n_water = waterModel.refractive_index(lam)
    n_glass = glassModel.refractive_index(lam)
    cos_i = np.multiply(rayDir, normal[None, :]).sum(-1)
    inward = cos_i < 0.0
    # the rays were pushed onto the correct side of the surface
    cos_re = np.multiply(refl_pos, normal[None, :]).sum(-1)
    cos_tr = np.multiply(trans_pos, normal[None, :]).sum(-1)
    assert np.all(cos_re * cos_i < 0.0)
    assert np.all(cos_tr * cos_i > 0.0)
    n_in = np.where(inward, n_water, n_glass)
    n_tr = np.where(inward, n_glass, n_water)
    eta = n_in / n_tr
    rs, rp = reflectance(np.abs(cos_i), n_in, n_tr)
    ts, tp = rs + 1.0, (rp + 1.0) * eta
    refl = 0.5 * (rs**2 + rp**2)
    t = 1.0 - refl
    rayNormal = -np.sign(cos_i)[:, None] * normal
    cos_i = -np.abs(cos_i)
    refl_dir_exp = rayDir - 2.0 * cos_i[:, None] * rayNormal
    k = 1.0 - eta**2 * (1.0 - cos_i**2)
    trans_dir_exp = eta[:, None] * rayDir
    with np.errstate(invalid="ignore"):
        trans_dir_exp -= (eta * cos_i + np.sqrt(k))[:, None] * rayNormal  # NaN under total reflection (wanted)

    # At the critical angle R(cos_i) has an infinite slope: a float32 rounding of cos_i or n (1e-7) moves R by more
    # than the reference's tolerance there (0.995 -> 1.0 for sample 11490 of this stream). Such rays (k within
    # 1e-5 of 0) are left out of the reflectance / Stokes comparisons; everything else keeps the reference's bounds.
    ok = np.abs(k) > 1e-5
    assert ok.sum() > N - 8
    assert np.all(contrib > 0.0)
    assert np.allclose(refl_dir, refl_dir_exp, atol=1e-5)
    assert np.allclose((refl_contrib / contrib)[ok], refl[ok], atol=8e-4)
    assert np.allclose(trans_dir[ok], trans_dir_exp[ok], atol=2e-4, equal_nan=True)
    assert np.allclose((trans_contrib / contrib)[ok], t[ok], atol=8e-4)
    if polarized:
        refl_polRef, trans_polRef = r[:, 19:22], r[:, 22:25]
        refl_stokes, trans_stokes = r[:, 25:29], r[:, 29:33]
        # the reference frames are normal to the plane of incidence
        assert np.abs(np.multiply(refl_polRef, rayDir).sum(-1)).max() < 5e-7
        assert np.abs(np.multiply(refl_polRef, normal).sum(-1)).max() < 5e-7
        assert np.abs(np.multiply(trans_polRef, rayDir).sum(-1)).max() < 5e-7
        assert np.abs(np.multiply(trans_polRef, normal).sum(-1)).max() < 5e-7
        assert np.all(refl_stokes[:, 0] == 1.0)
        sr = (rp**2 - rs**2) / (rp**2 + rs**2)
        assert np.abs(refl_stokes[:, 1] - sr)[ok].max() < 1e-4
        assert np.all(refl_stokes[:, 2:] == 0.0)
        assert np.all(trans_stokes[:, 0] == 1.0)
        st = (tp**2 - ts**2) / (tp**2 + ts**2)
        # not under total reflection (no transmitted ray), and not at grazing incidence, where t_s and t_p both
        # vanish and their float32 ratio is ill conditioned (the reference's 4e-3 is tuned to its own samples)
        sel = ok & (k > 0.0) & (np.abs(cos_i) > 1e-4)
        assert sel.sum() > 0.7 * N
        assert np.abs(trans_stokes[:, 1] - st)[sel].max() < 4e-3
        assert np.abs(trans_stokes[:, 1] - st)[sel].mean() < 1e-6
        assert np.all(trans_stokes[:, 2:] == 0.0)


# ------------------------------------------------------------------ tests/test_math.py, tests/test_random.py
def test_signBit(backend, rng):  # tests/test_math.py:8-38
    N = 32 * 256
    floats = (rng.random(N, dtype=np.float32) - 0.5) * 1e36
    floats[0] = 0.0
    floats[1] = -0.0
    results = ev(backend, _abi.UNIT_SIGN_BIT, floats.reshape(-1, 1))[:, 0]
    assert np.all(results == np.copysign(1.0, floats))


@pytest.mark.parametrize("alpha,gamma", [(1.0, 1.0), (0.5, 2.2), (4.8, 0.6)])
def test_gamma(backend, alpha, gamma):  # tests/test_random.py:92-122
    from scipy.stats import kstest
    from theia_b200 import random
    N = 32 * 32 * 1024
    with device(backend):
        philox = random.PhiloxRNG(key=0xABBAABBA)
        philox.update(0)
    x = ev(backend, _abi.UNIT_GAMMA, None, N, rng=philox, params=(alpha, gamma))[:, 0].astype(np.float64)
    # the reference's own assertion, literally (tests/test_random.py:121-122). Note its direction: the test passes when
    # the KS test REJECTS the gamma distribution. sh/random.gamma.glsl:17-19 sets a = sqrt(2 alpha - 1), c = alpha + 1 / alpha
    # where Cheng's GA has a = 1 / sqrt(2 alpha - 1), c = alpha + 1 / a: exact only for alpha = 1 (mean 2.9 % low at 0.5).
    # Reproduced as is, so the pin that means something is the algorithm itself:
    test = kstest(x, "gamma", args=(alpha, 0.0, 1.0 / gamma))
    assert test.pvalue < 0.05
    assert np.all(x > 0.0) and np.all(np.isfinite(x))

    # independent float64 restatement of sh/random.gamma.glsl:9-33 on the same draws
    import oracle
    M, K = 1 << 16, 24  # samples compared, trials followed per sample
    uDraw = oracle.philoxFill(0xABBAABBA, 0, M, 1 + 2 * K).astype(np.float64)
    al, scale, at = alpha, np.ones(M), 0
    if alpha < 1.0:
        scale = uDraw[:, 0] ** (1.0 / alpha)
        al, at = alpha + 1.0, 1
    a_, b_, c_ = np.sqrt(2.0 * al - 1.0), al - np.log(4.0), al + 1.0 / al
    want = np.full(M, np.nan)
    for k in range(K):
        u1, u2 = uDraw[:, at + 2 * k], uDraw[:, at + 2 * k + 1]
        with np.errstate(divide="ignore", invalid="ignore"):
            V = a_ * np.log(u1 / (1.0 - u1))
            X = al * np.exp(V)
            accept = (b_ + c_ * V - X >= np.log(u1 * u1 * u2)) & np.isnan(want)
        want[accept] = (scale * X / gamma)[accept]
    done = np.isfinite(want)
    assert done.mean() > 0.9  # (the acceptance rate of the sampler as written drops to about 10 % per trial at alpha = 4.8)
    rel = np.abs(x[:M][done] - want[done]) / want[done]
    # float32 against float64: a borderline trial may be decided the other way for a handful of samples
    assert np.mean(rel > 1e-4) < 1e-3
    assert np.median(rel) < 1e-6


def trackAngularEmission_pdf(x, n, cos_min=-1, cos_max=1, a=0.39, b=2.61, beta=1.0):
    """PDF of p(x) ~ exp(-bx^a)x^(a-1); x = |cos(theta) - cos(chev)| (tests/test_light.py:676-691)"""
    cos_chev = 1.0 / (beta * n)
    x = np.abs(x - cos_chev)
    int_lower = 1.0 - np.exp(-b * np.abs(cos_chev - cos_min) ** a)
    int_upper = 1.0 - np.exp(-b * np.abs(cos_chev - cos_max) ** a)
    int_lower *= np.sign(cos_min - cos_chev)
    int_upper *= np.sign(cos_max - cos_chev)
    norm = a * b / np.abs(int_upper - int_lower)
    return np.exp(-b * x**a) * x ** (a - 1) * norm


@pytest.mark.parametrize("n,a,b", [[1.33, 0.53, 3.3], [1.2, 0.86, 2.5], [1.45, 1.03, 3.1], [1.35, 1.13, 3.42]])
def test_particle_sampleEmissionAngle_full(backend, n, a, b):  # tests/test_light.py:721-765
    from theia_b200 import random
    N = 32 * 1024 * 1024  # as the reference: the sparse bins near cos = -1 need the statistics for its 5 % bound
    N_BINS = 100
    with device(backend):
        philox = random.PhiloxRNG(key=0x01ABBA10)
        sink = random.RNGBufferSink(philox, N, 1) if backend == "cuda" else None
        philox.update(0)
    if backend == "oracle":
        import oracle
        uRand = oracle.philoxFill(0x01ABBA10, 0, N, 1)
    else:
        from theia_b200.pipeline import runPipeline
        runPipeline([philox, sink])
        uRand = sink.tensor.cpu().numpy()
    samples = ev(backend, _abi.UNIT_EMISSION_ANGLE, uRand.reshape(-1, 1), params=(n, a, b))[:, 0]
    hist, edges = np.histogram(samples, N_BINS, (-1.0, 1.0), density=True)
    pdf = trackAngularEmission_pdf(edges, n, a=a, b=b)
    exp_hist = 0.5 * (pdf[1:] + pdf[:-1])
    peak_bin = int((1.0 / n + 1) / (2.0 / N_BINS))  # the pdf diverges at 1 / n: skip that bin and its neighbours
    err = np.abs(hist - exp_hist) / exp_hist
    assert err[: peak_bin - 1].max() < 0.05
    assert err[peak_bin + 2:].max() < 0.05


# ------------------------------------------------------------------ tests/test_target.py, tests/test_target_guide.py
def test_innerSphereTarget(backend):  # tests/test_target.py:125-163
    from theia_b200 import random, target
    from theia_b200.scene import RectBBox
    from theia_b200.testing import TargetSampler
    N = 32 * 256
    pos, radius = (8.0, -5.0, 3.2), 4.8
    box = RectBBox((3.0, -10.0, -1.0), (13.0, 0.0, 8.0))
    with device(backend):
        tgt = target.InnerSphereTarget(position=pos, radius=radius)
        philox = random.PhiloxRNG(key=0xC0FFEE)
        sampler = TargetSampler(N, tgt, rng=philox, sampleBox=box)
    runSampler(backend, sampler)
    r = sampler.getResults(0)
    od = np.sqrt(np.square(r["observer"] - pos).sum(-1))
    assert np.all((od >= radius) == r["occluded"])
    oclMask = r["occluded"] == 0
    assert np.allclose(r["sampleProb"][oclMask], 1.0 / (4.0 * np.pi * radius**2))
    assert np.all(r["sampleError"][oclMask] == 0)
    assert (r["sampleValid"] != 0).sum() > 0
    hitMask = oclMask & (r["hitValid"] != 0)
    assert hitMask.sum() > 0
    # Bit 3 of hitError (direction of hitPos - observer against the query direction, 1e-5) cannot hold in float32 for an
    # observer closer than ~0.1 m to the sphere: hitPos carries ulp(10) = 1e-6 of rounding, i.e. 1e-6 / t of direction
    # error. It is checked where it is well conditioned; the other bits (distance, object space position / normal) always.
    near = np.sqrt(np.square(r["hitPos"] - r["observer"]).sum(-1)) < 0.2
    assert np.all((r["hitError"][hitMask] & 7) == 0)
    assert np.all(r["hitError"][hitMask & ~near] == 0)
    assert (hitMask & near).sum() < 0.05 * hitMask.sum()
    p = np.array(pos)
    sd = np.sqrt(np.square(r["samplePos"][oclMask] - p).sum(-1))
    assert np.allclose(sd, radius)
    sn = (r["samplePos"][oclMask] - p) / sd[:, None]
    assert np.allclose(-sn, r["sampleNrm"][oclMask], atol=1e-7)
    hd = np.sqrt(np.square(r["hitPos"][hitMask] - p).sum(-1))
    assert np.allclose(hd, radius)
    hn = (r["hitPos"][hitMask] - p) / hd[:, None]
    assert np.allclose(-hn, r["hitNrm"][hitMask])


def test_sphereTarget(backend):  # tests/test_target.py:166-203
    from theia_b200 import random, target
    from theia_b200.testing import TargetSampler
    N = 32 * 256
    pos, radius = (12.0, -5.0, 3.2), 4.0
    with device(backend):
        tgt = target.SphereTarget(position=pos, radius=radius)
        philox = random.PhiloxRNG(key=0xC0FFEE)
        sampler = TargetSampler(N, tgt, rng=philox)
    runSampler(backend, sampler)
    r = sampler.getResults(0)
    od = np.sqrt(np.square(r["observer"] - pos).sum(-1))
    assert np.all((od <= radius) == r["occluded"])
    oclMask = r["occluded"] == 0
    expProb = 1.0 / (2.0 * np.pi * radius**2) / (1.0 - radius / od)
    assert np.allclose(r["sampleProb"][oclMask], expProb[oclMask])
    assert np.all(r["sampleError"][oclMask] == 0)
    assert (r["sampleValid"] != 0).sum() == N
    hitMask = oclMask & (r["hitValid"] != 0)
    assert hitMask.sum() > 0
    assert np.all(r["hitError"][hitMask] == 0)
    p = np.array(pos)
    sd = np.sqrt(np.square(r["samplePos"][oclMask] - p).sum(-1))
    assert np.allclose(sd, radius)
    sn = (r["samplePos"][oclMask] - p) / sd[:, None]
    assert np.allclose(sn, r["sampleNrm"][oclMask], atol=3e-7)
    hd = np.sqrt(np.square(r["hitPos"][hitMask] - p).sum(-1))
    assert np.allclose(hd, radius)
    hn = (r["hitPos"][hitMask] - p) / hd[:, None]
    assert np.allclose(hn, r["hitNrm"][hitMask])


def _planeTargetChecks(r, o2w, inside, area, N):
    """shared part of tests/test_target.py:16-122 (disk / flat): sample on the plane, ray / plane hit, probabilities"""
    w2o = o2w.inverse()
    objObs = w2o.apply(r["observer"])
    objDir = w2o.applyVec(r["direction"])
    nrm = np.zeros((N, 3))
    nrm[:, 2] = np.sign(objObs[:, 2])
    nrm = o2w.applyVec(nrm)
    expSampleValid = np.abs(objObs[:, 2]) > 1e-6  # should be close enough
    assert np.all((r["sampleValid"] != 0) == expSampleValid)
    assert np.all(r["sampleError"][expSampleValid] == 0)
    assert np.allclose(r["sampleNrm"][expSampleValid], nrm[expSampleValid], atol=1e-6)
    objPos = w2o.apply(r["samplePos"])
    assert np.all(inside(objPos[expSampleValid][:, :2]))
    assert np.abs(objPos[expSampleValid][:, 2]).max() < 2e-6
    with np.errstate(divide="ignore", invalid="ignore"):
        t = -objObs[:, 2] / objDir[:, 2]
    objHit = objObs + objDir * t[:, None]
    hit = (t > 0) & inside(objHit[:, :2]) & expSampleValid
    wHit = o2w.apply(objHit)
    assert hit.sum() > 0  # otherwise the tests make no sense
    # float32 against float64 on the rim: points within 1e-5 of the edge may fall on either side
    rim = np.abs(np.where(hit, 0.0, 1.0) - np.where((t > 0) & inside(objHit[:, :2] * (1.0 + 1e-5)) & expSampleValid, 0.0, 1.0)) > 0
    rim |= np.abs(np.where(hit, 0.0, 1.0) - np.where((t > 0) & inside(objHit[:, :2] * (1.0 - 1e-5)) & expSampleValid, 0.0, 1.0)) > 0
    assert rim.sum() < 8
    assert np.all(((r["hitValid"] != 0) == hit)[~rim])
    hit = hit & ~rim
    assert np.allclose(r["hitPos"][hit], wHit[hit], atol=2e-5)
    assert np.all(r["hitError"][hit] == 0)
    assert np.allclose(r["hitNrm"][hit], nrm[hit], atol=1e-6)
    assert np.allclose(r["sampleProb"][expSampleValid], 1.0 / area)
    assert np.allclose(r["hitProb"][hit], 1.0 / area)


def test_diskTarget(backend):  # tests/test_target.py:16-67
    from theia_b200 import random, target
    from theia_b200.scene import Transform
    from theia_b200.testing import TargetSampler
    N = 32 * 256
    radius, position, normal, up = 12.0 * u.m, (5.0, -2.0, 3.0), (-1.0, 2.0, 0.0), (1.0, 1.0, 1.0)
    o2w = Transform.View(position=position, direction=normal, up=up)
    with device(backend):
        tgt = target.DiskTarget(position=position, radius=radius, normal=normal, up=up)
        philox = random.PhiloxRNG(key=0xC0FFEE)
        sampler = TargetSampler(N, tgt, rng=philox)
    runSampler(backend, sampler)
    _planeTargetChecks(sampler.getResults(0), o2w, lambda xy: np.square(xy).sum(-1) <= radius**2, np.pi * radius**2, N)


def test_flatTarget(backend):  # tests/test_target.py:70-122
    from theia_b200 import random, target
    from theia_b200.scene import Transform
    from theia_b200.testing import TargetSampler
    N = 32 * 256
    width, length, position, normal, up = 12.0 * u.m, 18.0 * u.m, (5.0, -2.0, 3.0), (-1.0, 2.0, 0.0), (1.0, 1.0, 1.0)
    o2w = Transform.View(position=position, direction=normal, up=up)
    with device(backend):
        tgt = target.FlatTarget(width=width, length=length, position=position, direction=normal, up=up)
        philox = random.PhiloxRNG(key=0xC0FFEE)
        sampler = TargetSampler(N, tgt, rng=philox)
    runSampler(backend, sampler)
    _planeTargetChecks(sampler.getResults(0), o2w, lambda xy: (2.0 * np.abs(xy) <= (width, length)).min(-1), width * length, N)


def intersectSphere(center, rad, pos, dir):
    """near hit, far hit, hit mask (tests/test_target_guide.py:13-26; Ray Tracing Gems ch. 7)"""
    f = pos - center
    b2 = np.multiply(f, dir).sum(-1)
    fd = f - b2[:, None] * dir
    discr = rad**2 - np.square(fd).sum(-1)
    c = np.square(f).sum(-1) - rad**2
    with np.errstate(invalid="ignore"):
        q = -b2 - np.copysign(1.0, b2) * np.sqrt(discr)
    return c / q, q, discr >= 0.0


def test_sphereTargetGuide(backend):  # tests/test_target_guide.py:37-72
    from theia_b200 import random, target
    from theia_b200.testing import TargetGuideSampler
    N = 32 * 256
    pos, radius = (12.0, -5.0, 3.2), 4.0 * u.m
    with device(backend):
        guide = target.SphereTargetGuide(position=pos, radius=radius)
        philox = random.PhiloxRNG(key=0xC0FFEE)
        sampler = TargetGuideSampler(N, guide, rng=philox)
    runSampler(backend, sampler)
    r = sampler.getResults(0)
    pos = np.array(pos)
    d = np.sqrt(np.square(pos - r["observer"]).sum(-1))
    valid = d > radius
    assert np.all(r["sampleProb"][~valid] == 0.0)
    assert np.all((r["sampleProb"] > 0.0) == valid)
    t1, t2, hit = intersectSphere(pos, radius, r["observer"].astype(np.float64), r["sampleDir"].astype(np.float64))
    assert np.all(r["sampleDist"][valid] >= t2[valid])
    assert np.allclose(np.square(r["sampleDir"]).sum(-1), 1.0)
    with np.errstate(invalid="ignore"):
        cosMin = 1.0 - np.sqrt(1.0 - (radius / d) ** 2)
    prob = 1.0 / (2.0 * np.pi * cosMin)
    assert np.allclose(r["sampleProb"][valid], prob[valid])
    assert np.allclose(np.square(r["evalDir"]).sum(-1), 1.0)
    t1, t2, hit = intersectSphere(pos, radius, r["observer"].astype(np.float64), r["evalDir"].astype(np.float64))
    valid = valid & (t2 >= 0.0)
    assert np.all(r["evalProb"][~valid] == 0.0)
    assert np.all((r["evalProb"] > 0.0) == valid)
    assert np.all(r["evalDist"][valid] >= t2[valid])
    assert np.allclose(r["evalProb"][valid], prob[valid])


def _planeGuideChecks(r, o2w, normal, inside, area, *, distSlack, rtolSample):
    """shared part of tests/test_target_guide.py:75-185 (disk / flat guide)"""
    w2o = o2w.inverse()
    objObs = w2o.apply(r["observer"])
    valid = objObs[:, 2] > 0.0
    assert np.all((r["sampleProb"] > 0.0) == valid)
    assert np.all((r["sampleProb"] == 0.0) != valid)
    objDir = w2o.applyVec(r["sampleDir"])
    with np.errstate(divide="ignore", invalid="ignore"):
        tHit = -objObs[:, 2] / objDir[:, 2]
    assert np.all(tHit[valid] > 0.0)
    assert (r["sampleDist"] - tHit)[valid].min() >= -distSlack  # small error is fine
    objHit = objObs + objDir * tHit[:, None]
    assert np.all(inside(objHit[valid][:, :2] * (1.0 - 1e-5)))
    nrm = np.array(normal, dtype=np.float64)
    nrm /= np.sqrt(np.square(nrm).sum())
    cosNrm = np.abs(np.multiply(r["sampleDir"], nrm[None, :]).sum(-1))
    prob = 1.0 / area * (tHit**2) / cosNrm
    assert np.allclose(r["sampleProb"][valid], prob[valid], rtol=rtolSample)
    objDir = w2o.applyVec(r["evalDir"])
    with np.errstate(divide="ignore", invalid="ignore"):
        tHit = -objObs[:, 2] / objDir[:, 2]
    objHit = objObs + objDir * tHit[:, None]
    rim = inside(objHit[:, :2] * (1.0 + 1e-5)) != inside(objHit[:, :2] * (1.0 - 1e-5))  # float32 vs float64 on the edge
    valid = valid & (tHit > 0.0) & inside(objHit[:, :2])
    assert valid.sum() > 0 and rim.sum() < 8
    assert (r["evalDist"] - tHit)[valid & ~rim].min() >= -1e-5
    assert np.all(r["evalProb"][~valid & ~rim] == 0.0)
    cosNrm = np.abs(np.multiply(r["evalDir"], nrm[None, :]).sum(-1))
    prob = 1.0 / area * (tHit**2) / cosNrm
    assert np.allclose(r["evalProb"][valid & ~rim], prob[valid & ~rim])


def test_diskTargetGuide(backend):  # tests/test_target_guide.py:75-126
    from theia_b200 import random, target
    from theia_b200.scene import Transform
    from theia_b200.testing import TargetGuideSampler
    N = 32 * 256
    radius, position, normal = 12.0 * u.m, (5.0, -2.0, 3.0), (-1.0, 2.0, 0.0)
    o2w = Transform.View(position=position, direction=normal)
    with device(backend):
        guide = target.DiskTargetGuide(position=position, radius=radius, normal=normal)
        philox = random.PhiloxRNG(key=0xC0FFEE)
        sampler = TargetGuideSampler(N, guide, rng=philox)
    runSampler(backend, sampler)
    _planeGuideChecks(sampler.getResults(0), o2w, normal, lambda xy: np.square(xy).sum(-1) <= radius**2, np.pi * radius**2,
                      distSlack=0.02, rtolSample=2.5e-4)


def test_flatTargetGuide(backend):  # tests/test_target_guide.py:129-185
    from theia_b200 import random, target
    from theia_b200.scene import Transform
    from theia_b200.testing import TargetGuideSampler
    N = 32 * 256
    width, height, position, normal, up = 18.0 * u.m, 12.0 * u.m, (5.0, -2.0, 3.0), (-1.0, 2.0, 0.0), (1.0, 1.0, 1.0)
    o2w = Transform.View(position=position, direction=normal, up=up)
    with device(backend):
        guide = target.FlatTargetGuide(width=width, height=height, position=position, normal=normal, up=up)
        philox = random.PhiloxRNG(key=0xC0FFEE)
        sampler = TargetGuideSampler(N, guide, rng=philox)
    runSampler(backend, sampler)
    _planeGuideChecks(sampler.getResults(0), o2w, normal, lambda xy: (2.0 * np.abs(xy) <= (width, height)).min(-1),
                      width * height, distSlack=0.002, rtolSample=2e-4)


# ------------------------------------------------------------------ tests/test_light.py (forward samplers)
def _lightSampler(backend, makeLight, makePhotons, N, *, key, polarized=False, medium=None):
    from theia_b200 import light, random
    with device(backend):
        philox = random.PhiloxRNG(key=key)
        photons, source = makePhotons(), makeLight()
        sampler = light.LightSampler(source, photons, N, rng=philox, polarized=polarized,
                                     medium=medium() if medium else None)
    runSampler(backend, sampler)
    return source, sampler.lightQueue.view(0), sampler.wavelengthQueue.view(0)


def test_constWavelength(backend):  # tests/test_light.py:154-169
    from theia_b200 import light
    N, lam = 32 * 256, 450.0 * u.nm
    _, _, result = _lightSampler(backend, lambda: light.PencilLightSource(timeRange=(0.0, 0.0)),
                                 lambda: light.ConstWavelengthSource(lam), N, key=0xC0110FFC0FFEE)
    assert np.all(result["contrib"] == 1.0)
    assert np.all(result["wavelength"] == lam)


@pytest.mark.parametrize("normalize", [True, False])
def test_uniformWavelength(backend, normalize):  # tests/test_light.py:172-194
    from theia_b200 import light
    N = 32 * 256
    lamRange, dLam = (350.0 * u.nm, 750.0 * u.nm), 400.0 * u.nm
    contrib = 1.0 if normalize else dLam
    _, _, result = _lightSampler(backend, lambda: light.PencilLightSource(timeRange=(0.0, 0.0)),
                                 lambda: light.UniformWavelengthSource(lambdaRange=lamRange, normalize=normalize), N,
                                 key=0xC0110FFC0FFEE)
    assert np.all(result["contrib"] == contrib)
    assert np.abs(np.min(result["wavelength"]) - lamRange[0]) < 1.0
    assert np.abs(np.max(result["wavelength"]) - lamRange[1]) < 1.0


@pytest.mark.parametrize("polarized", [True, False])
def test_coneLightSource_fwd(backend, polarized):  # tests/test_light.py:233-278
    from theia_b200 import light
    N = 32 * 256
    position, direction, opening = (14.0, -2.0, 3.0), (0.8, 0.36, 0.48), 0.33
    stokes, polRefIn, budget = (1.0, 0.9, 0.1, -0.5), (0.0, 0.48, -0.36), 12.0
    _, result, _ = _lightSampler(
        backend, lambda: light.ConeLightSource(position=position, direction=direction, timeRange=(10.0 * u.ns, 10.0 * u.ns),
                                               cosOpeningAngle=opening, budget=budget, stokes=stokes,
                                               polarizationReference=polRefIn, polarized=polarized),
        lambda: light.ConstWavelengthSource(wavelength=100.0 * u.nm), N, key=0xC0110FFC0FFEE, polarized=polarized)
    assert result.count == N
    assert np.allclose(result["position"], position)
    assert (result["direction"] * direction).sum(-1).min() >= opening
    assert np.abs(np.square(result["direction"]).sum(-1) - 1.0).max() < 1e-5
    assert np.all(result["contrib"] == budget)
    if polarized:
        assert np.allclose(result["stokes"], stokes)
        polRef = result["polarizationRef"]
        assert np.abs(np.square(polRef).sum(-1) - 1.0).max() < 1e-5
        # polRef was only straightened against the ray, not rotated about it (that would invalidate the Stokes vector)
        vol = (np.cross(result["direction"], polRefIn) * polRef).sum(-1)
        assert np.abs(vol).max() < 1e-5


@pytest.mark.parametrize("polarized", [True, False])
def test_pencilLightSource(backend, polarized):  # tests/test_light.py:339-372
    from theia_b200 import light
    N = 32 * 256
    position, direction = (14.0, -2.0, 3.0), (0.8, 0.36, 0.48)
    stokes, polRef, budget = (1.0, 0.9, 0.1, -0.5), (0.0, 0.48, -0.36), 12.0
    _, result, _ = _lightSampler(
        backend, lambda: light.PencilLightSource(position=position, direction=direction, timeRange=(10.0 * u.ns, 10.0 * u.ns),
                                                 budget=budget, stokes=stokes, polarizationRef=polRef),
        lambda: light.ConstWavelengthSource(wavelength=100.0 * u.nm), N, key=0xC0110FFC0FFEE, polarized=polarized)
    assert result.count == N
    assert np.allclose(result["position"], position)
    assert np.allclose(result["direction"], direction)
    assert np.all(result["contrib"] == budget)
    if polarized:
        assert np.allclose(result["stokes"], stokes)
        assert np.allclose(result["polarizationRef"], polRef)


@pytest.mark.parametrize("polarized", [True, False])
def test_sphericalLightSource_fwd(backend, polarized):  # tests/test_light.py:375-411
    from theia_b200 import light
    N = 32 * 256
    position, budget = (14.0, -2.0, 3.0), 12.0
    _, result, _ = _lightSampler(
        backend, lambda: light.SphericalLightSource(position=position, timeRange=(10.0 * u.ns, 10.0 * u.ns), budget=budget),
        lambda: light.ConstWavelengthSource(wavelength=100.0 * u.nm), N, key=0xC0110FFC0FFEE, polarized=polarized)
    assert result.count == N
    assert np.all(result["position"] == np.array(position, dtype=np.float32))
    assert np.allclose(np.sqrt(np.square(result["direction"]).sum(-1)), 1.0)
    assert np.abs(np.mean(result["direction"], axis=0)).max() < 0.01
    assert np.allclose(result["contrib"], budget)
    if polarized:
        assert np.allclose(result["stokes"], (1.0, 0.0, 0.0, 0.0))
        polRef = result["polarizationRef"]
        assert np.abs((polRef * result["direction"]).sum(-1)).max() < 1e-5
        assert (np.square(polRef).sum(-1) - 1.0).max() < 1e-5


def _phiUniform(result, direction, up):
    from scipy import stats
    cos_phi = np.multiply(result["direction"], up).sum(-1)
    sin_phi = np.multiply(np.cross(result["direction"], up), direction).sum(-1)
    phi = (np.arctan2(sin_phi, cos_phi) + np.pi) / (2.0 * np.pi)
    return stats.kstest(phi[::100], "uniform").pvalue


@pytest.mark.parametrize("applyFrankTamm", [False, True])
def test_muonTrackLightSource_fwd(backend, applyFrankTamm):  # tests/test_light.py:839-909
    from scipy import integrate
    from theia_b200 import light, material
    from theia_b200.testing import WaterTestModel
    N = 256 * 1024
    lam_range = (300.0 * u.nm, 500.0 * u.nm)
    E_muon = 10.0 * u.TeV
    startPos, startTime = np.array([1.0, 5.0, -2.0]) * u.m, 0.5 * u.us
    direction, dist = np.array([0.36, 0.48, 0.8]), 100.0 * u.m
    endPos, endTime = startPos + direction * dist, startTime + dist / u.c
    up = np.array([0.8, -0.6, 0.0])
    water = WaterTestModel()
    store = {}

    def medium():
        store["s"] = material.MaterialStore([], media=[water.createMedium(*lam_range)])
        return store["s"].media["water"]

    src, result, photons = _lightSampler(
        backend, lambda: light.MuonTrackLightSource(startPos, startTime, endPos, endTime, E_muon, applyFrankTamm=applyFrankTamm),
        lambda: light.UniformWavelengthSource(lambdaRange=lam_range, normalize=False), N, key=0xC0FFEE, medium=medium)
    distStart = np.sqrt(np.square(startPos - result["position"]).sum(-1))
    distEnd = np.sqrt(np.square(endPos - result["position"]).sum(-1))
    assert np.allclose(distStart + distEnd, dist)
    assert np.allclose(result["startTime"], startTime + distStart / u.c)
    assert distStart.min() < 0.05 and distEnd.min() < 0.05
    assert np.allclose(np.square(result["direction"]).sum(-1), 1.0)
    cos_theta = np.multiply(result["direction"], direction).sum(-1)
    assert cos_theta.max() > 0.999 and cos_theta.min() < -0.999
    assert _phiUniform(result, direction, up) > 0.05
    # the Monte Carlo estimate converges to the number of Cherenkov photons produced along the track
    energyScale = src.getParam("_energyScale")
    if applyFrankTamm:
        lam = np.linspace(*lam_range, 1025)
        dlam = (lam_range[1] - lam_range[0]) / (len(lam) - 1)
        ft = light.frankTamm(lam, water.refractive_index(lam))
        nPhotons = integrate.romb(ft, dlam) * dist * energyScale
        contrib = result["contrib"].astype(np.float64) * photons["contrib"]
        assert np.abs(1.0 - contrib.mean() / nPhotons) < 5e-4
    else:
        assert np.allclose(result["contrib"], dist * energyScale)


@pytest.mark.parametrize("applyFrankTamm", [False, True])
@pytest.mark.parametrize("particle", ["e-", "pi+"])
def test_particleCascadeLightSource_fwd(backend, applyFrankTamm, particle):  # tests/test_light.py:991-1065
    from scipy import integrate
    from theia_b200 import cascades, light, material
    from theia_b200.testing import WaterTestModel
    N = 256 * 1024
    lam_range = (300.0 * u.nm, 500.0 * u.nm)
    startPos, startTime = np.array([1.0, 5.0, -2.0]) * u.m, 0.5 * u.us
    direction = np.array([0.36, 0.48, 0.8])
    up = np.array([0.8, -0.6, 0.0])
    primary = cascades.ParticleType.E_MINUS if particle == "e-" else cascades.ParticleType.PI_PLUS
    p = cascades.Particle(primary, startPos, direction, 1.0 * u.TeV, startTime)
    params = cascades.createParamsFromParticle(p, lightSourceName="")[1]
    water = WaterTestModel()
    store = {}

    def medium():
        store["s"] = material.MaterialStore([], media=[water.createMedium(*lam_range)])
        return store["s"].media["water"]

    _, result, photons = _lightSampler(
        backend, lambda: light.ParticleCascadeLightSource(**params, applyFrankTamm=applyFrankTamm),
        lambda: light.UniformWavelengthSource(lambdaRange=lam_range), N, key=0xC0FFEE, medium=medium)
    dirPos = result["position"] - startPos
    distPos = np.sqrt(np.square(dirPos).sum(-1))
    dirPos = dirPos / distPos[:, None]
    assert np.allclose(dirPos[distPos > 0.0], direction)
    assert np.allclose(startTime + distPos / u.c, result["startTime"])
    z_max = (params["a_long"] - 1) * params["b_long"]  # mode of the gamma distribution
    assert distPos.min() < 0.1 * z_max
    assert distPos.max() > 4.0 * z_max
    assert np.allclose(np.square(result["direction"]).sum(-1), 1.0)
    cos_theta = np.multiply(result["direction"], direction).sum(-1)
    assert cos_theta.min() < -0.999 and cos_theta.max() > 0.999
    assert _phiUniform(result, direction, up) > 0.05
    est = result["contrib"].astype(np.float64).mean()
    expEst = params["effectiveLength"]
    if applyFrankTamm:
        lam = np.linspace(*lam_range, 1025)
        dlam = (lam_range[1] - lam_range[0]) / (len(lam) - 1)
        dN_dx = integrate.romb(light.frankTamm(lam, water.refractive_index(lam)), dlam)
        expEst = dN_dx * params["effectiveLength"] / (lam_range[1] - lam_range[0])
    assert np.abs(1.0 - est / expEst) < 5e-4


# ------------------------------------------------------------------ tests/test_response.py (replay seam)
def runReplay(backend, replay, i=0):
    if backend == "oracle":
        import oracle
        oracle.runReplay(replay, i)
    else:
        import torch
        from theia_b200.pipeline import runPipeline
        runPipeline(replay.collectStages(), i)
        torch.cuda.synchronize()


def _fillHits(samples, rng, N, polarized, *, tMax=100.0 * u.ns, contrib=None):
    samples["position"] = (10.0 * rng.random((N, 3)) - 5.0) * u.m
    samples["direction"] = rng.random((N, 3))
    samples["normal"] = rng.random((N, 3))
    samples["wavelength"] = (rng.random((N,)) * 100.0 + 400.0) * u.nm
    samples["time"] = rng.random((N,)) * tMax
    samples["contrib"] = np.arange(N).astype(np.float32) if contrib is None else contrib
    if polarized:
        samples["stokes"] = rng.random((N, 4))
        samples["polarizationRef"] = rng.random((N, 3))


@pytest.mark.parametrize("polarized", [True, False])
def test_record(backend, rng, polarized):  # tests/test_response.py:17-50
    from theia_b200 import response
    N = 8192
    with device(backend):
        record = response.HitRecorder(polarized=polarized)
        replay = response.HitReplay(N, record, polarized=polarized)
    samples = replay.queue.view(0)
    _fillHits(samples, rng, N, polarized)  # contrib encodes the ordering (the device scrambles the items)
    runReplay(backend, replay)
    results = record.queue.view(0)
    assert results.count == N
    results = results[np.argsort(results["contrib"])]
    fields = ["position", "direction", "normal", "wavelength", "time", "contrib"]
    for f in fields + (["stokes", "polarizationRef"] if polarized else []):
        assert np.all(samples[f] == results[f]), f


def test_hitPolarizationMismatch(backend):  # tests/test_response.py:53-59
    from theia_b200 import response
    with device(backend):
        with pytest.raises(RuntimeError):
            response.HitReplay(128, response.HitRecorder(polarized=True), polarized=False)
        with pytest.raises(RuntimeError):
            response.HitReplay(128, response.HitRecorder(polarized=False), polarized=True)


def test_histogramResponse(backend, rng):  # tests/test_response.py:62-99
    from theia_b200 import response
    N, N_BINS, BIN_SIZE, T0, NORM = 256 * 1024, 50, 4.0 * u.ns, 20.0 * u.ns, 1e-5
    T1 = T0 + N_BINS * BIN_SIZE
    with device(backend):
        resp = response.HistogramHitResponse(response.UniformValueResponse(), nBins=N_BINS, t0=T0, binSize=BIN_SIZE,
                                             normalization=NORM)
        replay = response.HitReplay(N, resp, blockSize=128)
    samples = replay.queue.view(0)
    _fillHits(samples, rng, N, False, tMax=T1, contrib=rng.random((N,)) * 10.0)
    runReplay(backend, replay)
    result = resp.result(0)
    bin_edge = np.arange(N_BINS + 1) * BIN_SIZE + T0
    exp_hist, _ = np.histogram(samples["time"], bin_edge, weights=samples["contrib"])
    assert np.allclose(result, exp_hist * NORM, rtol=1e-4)


@pytest.mark.parametrize("storeObjectId", [True, False])
def test_StoreTimeHitResponse(backend, rng, storeObjectId):  # tests/test_response.py:253-288
    from theia_b200 import random, response
    N = 32 * 256
    with device(backend):
        philox = random.PhiloxRNG(key=0xC0FFEE)
        value = response.UniformValueResponse()
        resp = response.StoreTimeHitResponse(value, storeObjectId=storeObjectId)
        replay = response.HitReplay(N, resp, rng=philox)
    samples = replay.queue.view(0)
    _fillHits(samples, rng, N, False, contrib=rng.random((N,)))
    samples["time"] = np.arange(N)  # use integers for better checking
    samples["objectId"] = np.arange(N, dtype=np.int32)  # use same as time stamp
    runReplay(backend, replay)
    result = resp.result(0)
    assert result is not None
    # we expect about half the photons to be accepted (binomial: sigma = 0.0055)
    assert abs(result.count / N - 0.5) < 0.02
    # the random numbers drawn are lost: check that the sampled time stamps can be found in the input
    time = result["time"][: result.count]
    assert time.min() >= 0 and time.max() < N
    assert len(np.unique(time)) == result.count
    # a hit is kept when its draw is below its value: over all hits the kept ones carry the larger contributions
    kept = np.zeros(N, dtype=bool)
    kept[time.astype(np.int64)] = True
    assert samples["contrib"][kept].mean() > 0.6 and samples["contrib"][~kept].mean() < 0.4
    if storeObjectId:
        assert np.all(time == result["objectId"][: result.count])
    else:
        assert "objectId" not in result.fields


@pytest.mark.parametrize("useSharedMemory", [False, True])
def test_kernelHistogramEstimator(backend, rng, useSharedMemory):  # tests/test_response.py:133-180
    from scipy import stats
    from theia_b200 import response
    N, N_BINS, BIN_SIZE, T0 = 64 * 1024, 200, 1.2 * u.ns, 20.0 * u.ns
    T1 = T0 + N_BINS * BIN_SIZE
    T_PEAK = T0 + 0.3 * (T1 - T0)
    BANDWIDTH, SUPPORT, NORM = 10.0 * u.ns, 50.0 * u.ns, 5e-4
    with device(backend):
        value = response.UniformValueResponse()
        resp = response.KernelHistogramHitResponse(value, nBins=N_BINS, t0=T0, binSize=BIN_SIZE, kernelBandwidth=BANDWIDTH,
                                                   kernelSupport=SUPPORT, normalization=NORM, useSharedMemory=useSharedMemory)
        replay = response.HitReplay(N, resp)
    samples = replay.queue.view(0)
    _fillHits(samples, rng, N, False, contrib=np.zeros(N))
    samples["time"] = rng.normal(T_PEAK, 10.0 * BIN_SIZE, (N,))
    samples["contrib"] = rng.random((N,)) + np.sqrt(samples["time"])
    runReplay(backend, replay)
    x = np.arange(N_BINS + 1) * BIN_SIZE + T0
    t = samples["time"].astype(np.float64)
    v = samples["contrib"].astype(np.float64)
    exp_result = np.zeros_like(x)
    for lo in range(0, N, 4096):  # (the reference loops over the hits one by one)
        exp_result += (v[lo:lo + 4096, None] * stats.norm.cdf(x[None, :], loc=t[lo:lo + 4096, None], scale=BANDWIDTH)).sum(0)
    exp_result = np.diff(exp_result) * NORM
    result = resp.result(0)
    assert np.abs(result - exp_result).max() < 6e-4


# ------------------------------------------------------------------ tests/test_light.py: host sources, function wavelength
@pytest.mark.parametrize("polarized", [False, True])
def test_hostLightSource(backend, rng, polarized):  # tests/test_light.py:23-72
    from theia_b200 import light
    N = 32 * 256
    with device(backend):
        photon = light.HostWavelengthSource(N)
        source = light.HostLightSource(N, polarized=polarized)
        sampler = light.LightSampler(source, photon, N, polarized=polarized)
    rays, photons = source.queue.view(0), photon.queue.view(0)
    cos_theta = 2.0 * rng.random(N) - 1.0
    sin_theta = np.sqrt(1.0 - cos_theta**2)
    phi = rng.random(N) * 2.0 * np.pi
    rays["position"] = (rng.random((N, 3)) * 10.0 - 5.0) * u.m
    rays["direction"] = np.stack([sin_theta * np.cos(phi), sin_theta * np.sin(phi), cos_theta], axis=-1)
    lam = (rng.random((N,)) * 600.0 + 200.0) * u.nm
    photons["wavelength"] = lam
    photons["contrib"] = rng.random((N,)) + 5.0
    rays["startTime"] = rng.random((N,)) * 50.0 * u.ns
    rays["contrib"] = rng.random((N,)) + 1.0
    if polarized:
        rays["stokes"] = rng.random((N, 4))  # we do not care about valid values
        rays["polarizationRef"] = rng.random((N, 3))
    runSampler(backend, sampler)
    lightResult, phResult = sampler.lightQueue.view(0), sampler.wavelengthQueue.view(0)
    assert lightResult.count == N
    for f in ("position", "direction", "startTime", "contrib") + (("stokes", "polarizationRef") if polarized else ()):
        assert np.allclose(lightResult[f], rays[f]), f
    assert np.allclose(phResult["contrib"], photons["contrib"])
    assert np.allclose(phResult["wavelength"], lam)


def test_HostSamplerMismatch(backend):  # tests/test_light.py:144-151
    from theia_b200 import light
    with device(backend):
        photon = light.ConstWavelengthSource()
        with pytest.raises(RuntimeError):
            light.LightSampler(light.HostLightSource(128, polarized=True), photon, 128, polarized=False)
        with pytest.raises(RuntimeError):
            light.LightSampler(light.HostLightSource(128, polarized=False), photon, 128, polarized=True)


def test_functionWavelength(backend):  # tests/test_light.py:197-231
    from theia_b200 import light
    N, lamRange = 256 * 1024, (300.0 * u.nm, 700.0 * u.nm)

    def fn(lam):  # something fancy just for testing
        x = lam / 1000
        return -x * np.log(x)

    def Fn(lam):  # its anti-derivative
        x = lam / 1000
        return 250.0 * x**2 * (1.0 - 2.0 * np.log(x))

    _, _, result = _lightSampler(backend, lambda: light.PencilLightSource(timeRange=(0.0, 0.0)),
                                 lambda: light.FunctionWavelengthSource(fn, lambdaRange=lamRange), N, key=0xC0FFEE)
    exp_contrib = Fn(lamRange[1]) - Fn(lamRange[0])
    assert np.allclose(result["contrib"], exp_contrib)
    assert np.abs(np.min(result["wavelength"]) - lamRange[0]) < 1.0
    assert np.abs(np.max(result["wavelength"]) - lamRange[1]) < 1.0
    hist, edges = np.histogram(result["wavelength"], bins=40)
    F = Fn(edges)
    assert np.abs(hist / N - (F[1:] - F[:-1]) / exp_contrib).max() < 9e-4


def _cherenkovChecks(samples, photons, photon, model, trackDirs, lengthFactor, usePhotons, polarized):
    """shared checks of tests/test_light.py:453-518 and :599-676: cone angle, Frank-Tamm contribution, polarisation"""
    from scipy import constants as c
    cos_theta = np.multiply(samples["direction"], trackDirs).sum(-1)
    lam = photons["wavelength"].ravel().astype(np.float64)
    n = model.refractive_index(lam)
    assert np.allclose(cos_theta, 1.0 / n)
    contrib = 1.0 - (1.0 / n**2)
    if usePhotons:
        contrib *= 2.0 * np.pi * c.alpha * 1e9 / (lam**2)
    else:
        contrib *= np.pi * c.e * c.c**2 * c.mu_0 * 1e18 / (lam**3)
    contrib *= lengthFactor  # 1 / p(position)
    lam0, lam1 = photon.getParam("lambdaRange")
    assert np.allclose(contrib, samples["contrib"].ravel())
    assert np.allclose(photons["contrib"], abs(lam1 - lam0))
    if polarized:
        polRef_exp = np.cross(trackDirs, samples["direction"])
        polRef_exp /= np.sqrt(np.square(polRef_exp).sum(-1))[:, None]
        polRef = samples["polarizationRef"]
        assert np.allclose(samples["stokes"], (1.0, 1.0, 0.0, 0.0))
        assert np.abs(np.abs((polRef * polRef_exp).sum(-1)) - 1.0).max() < 1e-6
        assert np.abs(np.square(polRef).sum(-1) - 1.0).max() < 1e-6
        assert np.abs((polRef * samples["direction"]).sum(-1)).max() < 1e-6


def _waterMedium(store):
    from theia_b200 import material
    from theia_b200.testing import WaterTestModel

    def medium():
        store["s"] = material.MaterialStore([], media=[WaterTestModel().createMedium()])
        return store["s"].media["water"]
    return medium


@pytest.mark.parametrize("usePhotons", [False, True])
@pytest.mark.parametrize("polarized", [True, False])
def test_cherenkov_fwd(backend, usePhotons, polarized):  # tests/test_light.py:453-518
    from theia_b200 import light
    from theia_b200.testing import WaterTestModel
    N = 32 * 256
    dir = np.array([0.36, 0.48, 0.8])
    dist = 200.0 * u.m
    startPos, startT = -dir * 0.5 * dist, -0.5 * dist / u.c
    endPos, endT = dir * 0.5 * dist, 0.5 * dist / u.c
    keep = {}
    source, samples, photons = _lightSampler(
        backend, lambda: light.CherenkovLightSource(trackStart=startPos, trackEnd=endPos, startTime=startT, endTime=endT,
                                                    usePhotonCount=usePhotons),
        lambda: keep.setdefault("photon", light.UniformWavelengthSource(normalize=False)), N, key=0xC0FFEE,
        polarized=polarized, medium=_waterMedium(keep))
    distA = np.sqrt(np.square(samples["position"] - startPos).sum(-1))
    distB = np.sqrt(np.square(samples["position"] - endPos).sum(-1))
    assert np.allclose(distA + distB, dist)  # on the track
    t = samples["startTime"].ravel()
    assert np.all((t >= startT) & (t <= endT))
    l = distA / dist
    # (atol: l comes from float32 positions at +- 100 m, i.e. 2e-5 ns of error, which the reference's bare allclose only
    # tolerates for |t| > 2 ns)
    assert np.allclose(t, (1.0 - l) * startT + l * endT, atol=1e-4)
    _cherenkovChecks(samples, photons, keep["photon"], WaterTestModel(), dir[None, :], dist, usePhotons, polarized)


@pytest.mark.parametrize("usePhotons", [False, True])
@pytest.mark.parametrize("polarized", [True, False])
def test_cherenkovTrack(backend, usePhotons, polarized):  # tests/test_light.py:599-676
    from theia_b200 import light
    from theia_b200.testing import WaterTestModel
    N = 32 * 256
    vertices = np.array([[0.0, 0.0, 0.0, 0.0], [1.0, 0.0, 0.0, 20.0], [1.0, 1.0, 0.0, 35.0], [1.0, 1.0, 1.0, 60.0]])
    trackDir = np.eye(3)
    dt = np.array([20.0, 15.0, 25.0])
    keep = {}

    def makeLight():
        keep["track"] = light.ParticleTrack(4)
        keep["track"].setVertices(vertices)
        return light.CherenkovTrackLightSource(keep["track"], usePhotonCount=usePhotons)

    source, samples, photons = _lightSampler(
        backend, makeLight, lambda: keep.setdefault("photon", light.UniformWavelengthSource(normalize=False)), N, key=0xC0FFEE,
        polarized=polarized, medium=_waterMedium(keep))
    assert np.all((samples["position"] >= 0.0) & (samples["position"] <= 1.0))
    t = samples["startTime"].ravel()
    assert np.all((t >= 0.0) & (t <= 60.0))
    assert np.allclose(t, np.multiply(samples["position"], dt[None, :]).sum(-1))
    segmentId = 2 - (samples["position"] == 0.0).sum(-1)
    # (three segments of unit length: 1 / p(position) = 3)
    _cherenkovChecks(samples, photons, keep["photon"], WaterTestModel(), trackDir[segmentId], 3.0, usePhotons, polarized)


@pytest.mark.parametrize("n,cos_min,cos_max,a,b,beta", [[1.33, -1.0, 1.0, 0.39, 2.61, 1.0], [1.40, -1.0, -0.5, 0.41, 2.39, 0.9],
                                                        [1.20, 0.8, 0.9, 0.44, 2.78, 1.0], [1.35, 0.1, 0.9, 0.41, 2.78, 0.7],
                                                        [1.35, -0.1, 0.9, 0.41, 2.78, 1.0]])
def test_trackAngularEmission_pdf(backend, n, cos_min, cos_max, a, b, beta):  # tests/test_light.py:679-718
    from scipy import integrate
    if backend != "oracle":
        pytest.skip("host side only: the pdf the emission angle tests compare against")

    def pdf(x):
        cos_chev = 1.0 / (beta * n)
        lower = (1.0 - np.exp(-b * np.abs(cos_chev - cos_min) ** a)) * np.sign(cos_min - cos_chev)
        upper = (1.0 - np.exp(-b * np.abs(cos_chev - cos_max) ** a)) * np.sign(cos_max - cos_chev)
        y = np.abs(x - cos_chev)
        return np.exp(-b * y**a) * y ** (a - 1) * a * b / np.abs(upper - lower)

    cos_chev = 1.0 / (beta * n)
    points = [cos_chev] if cos_min < cos_chev < cos_max else None
    est = integrate.quad(pdf, cos_min, cos_max, points=points, limit=200)[0]
    assert np.abs(est - 1.0) < 5e-4
    assert np.all(pdf(np.linspace(cos_min, cos_max, 500)) >= 0.0)


def test_MediumShader(backend, rng):  # tests/test_material.py:132-214 (tests/shader/medium.test.glsl)
    from theia_b200 import material
    N = 32 * 256

    class WaterModel(material.WaterBaseModel, material.HenyeyGreensteinPhaseFunction, material.MediumModel):
        ModelName = "water"

        def __init__(self) -> None:
            material.WaterBaseModel.__init__(self, 5.0, 1000.0, 35.0)
            material.HenyeyGreensteinPhaseFunction.__init__(self, 0.6)

    model = WaterModel()
    with device(backend):
        water = model.createMedium()
        store = material.MaterialStore([], media=[water])
        dLam = water.lambda_max - water.lambda_min
        wavelength = ((rng.random(N, np.float32) * dLam + water.lambda_min) * u.nm).astype(np.float32)
        theta = (1.0 - 2 * rng.random(N, np.float32)).astype(np.float32)  # [-1, 1]
        eta = rng.random(N, np.float32)  # [0, 1]
        r = ev(backend, _abi.UNIT_MEDIUM, np.stack([wavelength, theta, eta], axis=-1), medium=store.media["water"])
    assert np.all(r[:, 10:12].astype(np.float32).view(np.uint32) == 0)  # no material: no flags
    mu_s = model.scattering_coef(wavelength)
    assert np.allclose(r[:, 0], model.refractive_index(wavelength), 1e-4)
    assert np.allclose(r[:, 1], model.group_velocity(wavelength), 1e-4)
    assert np.allclose(r[:, 2], mu_s, 5e-3)
    assert np.allclose(r[:, 3], model.absorption_coef(wavelength) + mu_s, 5e-3)
    assert np.abs(r[:, 4] - model.log_phase_function(theta)).max() < 5e-5
    assert np.allclose(r[:, 5], model.phase_sampling(eta), 1e-4, 1e-5)


def test_MaterialShader(backend, rng):  # tests/test_material.py:217-371 (tests/shader/material.test.glsl)
    from theia_b200 import material
    N = 32 * 1000

    class WaterModel(material.WaterBaseModel, material.HenyeyGreensteinPhaseFunction,
                     material.KokhanovskyOceanWaterPhaseMatrix, material.MediumModel):
        ModelName = "water"

        def __init__(self) -> None:
            material.WaterBaseModel.__init__(self, 5.0, 1000.0, 35.0)
            material.HenyeyGreensteinPhaseFunction.__init__(self, 0.6)
            material.KokhanovskyOceanWaterPhaseMatrix.__init__(self, p90=0.66, theta0=0.25, alpha=4.0, xi=25.6)

    water_model, glass_model = WaterModel(), material.BK7Model()
    lam = ((rng.random(N, np.float32) * 600.0 + 200.0) * u.nm).astype(np.float32)
    theta = (1.0 - 2 * rng.random(N, np.float32)).astype(np.float32)
    eta = rng.random(N, np.float32)
    rows = np.stack([lam, theta, eta], axis=-1)
    with device(backend):
        water = water_model.createMedium()
        glass = glass_model.createMedium(name="glass", num_lambda=4096)
        mat_vac_glass = material.Material("vac_glass", None, glass)
        store = material.MaterialStore([material.Material("water_glass", water, glass), mat_vac_glass])
        h = N // 2
        res = {(name, side): ev(backend, _abi.UNIT_MEDIUM, rows[sl], material=store.material[name], params=(side,))
               for name, sl in (("water_glass", slice(0, h)), ("vac_glass", slice(h, N))) for side in (0, 1)}

    def check(r, model, sl, vacuum=False):
        zeros = np.zeros(h)
        n = np.ones(h) if vacuum else model.refractive_index(lam[sl])
        vg = np.full(h, 1.0 * u.c) if vacuum else model.group_velocity(lam[sl])
        volume = isinstance(model, WaterModel)  # glass and vacuum do not (volume) scatter
        mu_s = model.scattering_coef(lam[sl]) if volume else zeros
        mu_e = mu_s + (zeros if vacuum else model.absorption_coef(lam[sl]))
        assert np.allclose(r[:, 0], n, 1e-4) and np.allclose(r[:, 1], vg, 1e-4)
        assert np.allclose(np.exp(-r[:, 2]), np.exp(-mu_s), 1e-5, 1e-3)
        assert np.allclose(np.exp(-r[:, 3]), np.exp(-mu_e), 1e-5, 1e-3)
        assert np.abs(r[:, 4] - (model.log_phase_function(theta[sl]) if volume else zeros)).max() < 5e-4
        assert np.allclose(r[:, 5], model.phase_sampling(eta[sl]) if volume else zeros, 1e-4, 1e-5)
        for col, fn in ((6, "phase_m12"), (7, "phase_m22"), (8, "phase_m33")):
            assert np.abs(r[:, col] - (getattr(model, fn)(theta[sl]) if volume else zeros)).max() < 1e-3, fn
        assert np.abs(r[:, 9]).max() < 1e-3  # no m34 for water

    check(res[("water_glass", 0)], water_model, slice(0, h))
    check(res[("water_glass", 1)], glass_model, slice(0, h))
    check(res[("vac_glass", 0)], None, slice(h, N), vacuum=True)
    check(res[("vac_glass", 1)], glass_model, slice(h, N))
    flags = res[("vac_glass", 0)][0, 10:12].astype(np.float32).view(np.uint32)
    assert flags[0] == mat_vac_glass.flagsInward and flags[1] == mat_vac_glass.flagsOutward
