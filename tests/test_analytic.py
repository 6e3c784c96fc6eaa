"""
Analytic acceptance tests ported from the reference's own test-suite (they need no reference
binary). Each runs on the CPU oracle (pinning the restatement; part of the "not gpu" suite,
at reduced statistics) and on the CUDA path (marked gpu, at the reference's statistics).
"""

import numpy as np
import pytest

import harness as H
import theia_b200.units as u
from theia_b200 import _device

DEVICES = [pytest.param("cpu", id="oracle"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("flag,reflectance,err", [("T", 0.0, 0.0), ("R", 1.0, 0.0), ("TR", 0.0516, 0.005)])
def test_beam_splitter_reflectance(device, flag, reflectance, err):
    # tests/test_trace.py:1196-1275: Fresnel reflectance of BK7 at 45 degrees, 500 nm
    N = 32 * 256
    counts = []
    with _device.useDevice(device):
        for targetId in (1, 2):
            cfg = H.beamSplitter(N, flag, targetId)
            H.runAny(cfg)
            counts.append(cfg["response"].queue.view(0).count)
    assert counts[0] + counts[1] == N
    assert abs(counts[0] / N - reflectance) <= err


ENERGY_ROWS = [
    (0.0, 0.005, 0.05, 0.0, False, True, False, 4.5e-4),
    (0.0, 0.005, 0.05, 0.0, False, False, False, 0.011),
    (0.0, 0.005, 0.05, 0.0, True, True, False, 0.011),
    (0.05, 0.01, 0.05, 0.0, False, True, False, 1.8e-3),
    (0.05, 0.01, 0.05, 0.0, True, True, False, 2.6e-3),
    (0.05, 0.01, 0.05, -0.9, False, True, False, 0.012),
    (0.05, 0.01, 0.05, 0.9, True, True, False, 2.8e-4),
    (0.05, 0.01, 0.05, 0.9, True, False, False, 2.9e-3),
    (0.05, 0.01, np.nan, 0.9, False, True, True, 2.9e-3),
    (0.0, 0.005, 0.05, 0.0, False, True, True, 5e-4),
    (0.0, 0.005, 0.05, 0.0, True, True, True, 1.2e-3),
    (0.05, 0.01, 0.05, -0.9, False, True, True, 0.012),
]


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("mu_a,mu_s,mu_sample,g,disableDirect,sampleTarget,polarized,err", ENERGY_ROWS)
def test_volume_tracer_energy_conservation(device, mu_a, mu_s, mu_sample, g, disableDirect, sampleTarget, polarized, err):
    # tests/test_trace_energy.py:816-938: point source inside an InnerSphereTarget; the
    # un-attenuated contributions must add up to the light budget
    if device == "cuda":
        N, nBatches, tol = 512 * 1024, 10, err
    else:
        N, nBatches = 256 * 1024, 4
        tol = max(3.0 * err * np.sqrt(5.0), 3e-3)  # 5 times fewer paths than the reference's test
    budget, radius, t0 = 1e9, 100.0, 10.0
    with _device.useDevice(device):
        cfg = H.innerSphereEnergy(N, mu_a, mu_s, mu_sample, g, disableDirect, sampleTarget, polarized)
        total, stokes = 0.0, []
        for b in range(nBatches):
            H.runAny(cfg, b % 2)
            v = H.hitView(cfg, b % 2)
            vg = 1.0 / cfg["model"].ng * u.c
            d = vg * (v["time"].astype(np.float64) - t0)
            total += float((v["contrib"].astype(np.float64) * np.exp(mu_a * d)).sum())
            if polarized:
                stokes.append(np.array(v["stokes"]))
    direct = budget * np.exp(-mu_s * radius)
    expected = budget - direct if disableDirect else budget
    estimate = total / (N * nBatches)
    print(f"rel.err {estimate / expected - 1.0:+.2e} (tol {tol:.1e})")
    assert abs(estimate / expected - 1.0) < tol
    if polarized:
        s = np.concatenate(stokes)
        assert np.abs(s[:, 0] - 1.0).max() < 1e-5
        assert s[:, 1:].max() <= 1.0 and s[:, 1:].min() >= -1.0
        assert np.all(np.square(s[:, 1:]).sum(-1) <= 1.0 + 1e-5)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("shape", ["disk", "flat"])
def test_plane_target_solid_angle(device, shape):
    # Flat / disk targets (sh/target.flat.glsl, sh/target.disk.glsl): the unscattered light an
    # isotropic point source on the axis of the target deposits on it is budget * Omega / 4 pi,
    # Omega = solid angle of the shape. Only first-segment hits are summed: behind them the
    # reference clamps every free path at the target's PLANE whether or not the shape was hit
    # (dist = min(hit.dist, dist) with hit.dist = t also for invalid samples,
    # sh/tracer.volume.forward.glsl:163-167 with sh/target.flat.glsl:52-73), which is reproduced
    # here as it is and has no closed form.
    from theia_b200 import light, material, random, response, target, trace
    h, N = 2.0, 256 * 1024
    with _device.useDevice(device):
        med = H.HomogenousModel(mu_a=0.0, mu_s=1e-4).createMedium()
        store = material.MaterialStore([], media=[med])
        if shape == "disk":
            R = 1.5
            tgt = target.DiskTarget(position=(1.0, 2.0, 3.0), radius=R, normal=(0.0, 0.6, 0.8))
            omega = 2.0 * np.pi * (1.0 - h / np.hypot(h, R))
        else:
            a, b = 1.0, 2.5
            tgt = target.FlatTarget(width=a, length=b, position=(1.0, 2.0, 3.0), direction=(0.0, 0.6, 0.8),
                                    up=(1.0, 0.0, 0.0))
            omega = 4.0 * np.arctan(a * b / (2.0 * h * np.sqrt(4.0 * h * h + a * a + b * b)))
        src = light.SphericalLightSource(position=(1.0, 2.0 + 0.6 * h, 3.0 + 0.8 * h), timeRange=(0.0, 0.0), budget=1.0)
        rec = response.HitRecorder()
        rec.enableTags()
        rng = random.PhiloxRNG(key=0xD15C)
        tr = trace.VolumeForwardTracer(N, src, tgt, light.ConstWavelengthSource(500.0), rec, rng,
                                       medium=store.media[med.name], nScattering=2, scatterCoefficient=1e-4,
                                       disableTargetSampling=True, maxTime=float("inf"))
        cfg = {"tracer": tr, "response": rec, "callback": None, "rng": rng, "store": store}
        H.runAny(cfg)
        hits = H.sortedHits(cfg)
    first = (hits["tag"] & 0xFFFF) == 0
    total = float(hits["contrib"][first].astype(np.float64).sum()) / N
    # hits are reported in the target's frame: on the plane, inside the shape, normal +z = source side
    pos, nrm = hits["position"][first], hits["normal"][first]
    assert np.abs(pos[:, 2]).max() < 1e-5 and np.all(nrm[:, 2] == 1.0)
    if shape == "disk":
        assert np.hypot(pos[:, 0], pos[:, 1]).max() <= R * (1.0 + 1e-6)
    else:
        assert np.abs(pos[:, 0]).max() <= 0.5 * a * (1.0 + 1e-6) and np.abs(pos[:, 1]).max() <= 0.5 * b * (1.0 + 1e-6)
    expected = omega / (4.0 * np.pi)
    tol = 4.0 * np.sqrt((1.0 - expected) / (expected * N)) + 3e-4
    print(f"{shape}: {total:.6f} vs {expected:.6f}")
    assert abs(total / expected - 1.0) < tol


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("builder,kw,budget", [(H.volumeScenario, dict(mis=False), 1e5),
                                               (H.sceneC2, dict(lightKind="cone", polarized=False), 1e3),
                                               (H.sceneC2, dict(lightKind="cone", polarized=True), 1e3)],
                         ids=["volume", "c2", "c2-polarized"])
def test_photon_tracer_estimates_the_forward_tracers_yield(device, builder, kw, budget):
    # VolumePhotonTracer / ScenePhotonTracer (sh/tracer.*.photon.*.glsl): absorption roulette instead of
    # weights. Summed over hits, both estimate the same detected fraction of the emitted light.
    N = 1 << 17 if device == "cuda" else 1 << 16
    with _device.useDevice(device):
        fwd = builder(N, resp="record", **kw)
        H.runAny(fwd)
        a = H.hitView(fwd)["contrib"].astype(np.float64) / budget
        pho = builder(N, resp="record", photon=(3, 4), **kw)
        H.runAny(pho)
        b = H.hitView(pho)["contrib"].astype(np.float64)
    assert len(b) > 200 and b.max() <= 1.0 and b.min() > 0.0
    ya, yb = a.sum() / N, b.sum() / N
    err = np.sqrt((np.square(a).sum() + np.square(b).sum())) / N
    print(f"forward {ya:.5f} photon {yb:.5f} +- {err:.5f}")
    assert abs(ya - yb) < 4.5 * err


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("kind", ["cherenkov", "cherenkovTrack"])
def test_cherenkov_light_yield_and_cone(device, kind):
    # sh/lightsource.cherenkov.{simple,track}.glsl: emission on the Cherenkov cone cos(theta) = 1 / n, and the
    # light reaching a sphere around the track adds up to the Frank-Tamm yield 2 pi alpha / lambda^2 (1 - 1/n^2) L
    from theia_b200 import light, material, random, response, target, trace
    N, n, lam, L = 1 << 16, 1.33, 400.0, 12.0
    with _device.useDevice(device):
        med = H.EnergyModel(0.0, 0.005, 0.0, n=n, ng=n).createMedium()
        store = material.MaterialStore([], media=[med])
        if kind == "cherenkov":
            src = light.CherenkovLightSource(trackStart=(1.0, 2.0, 3.0), trackEnd=(1.0, 2.0 + 0.6 * L, 3.0 + 0.8 * L),
                                             startTime=5.0, endTime=5.0 + L / 0.299792458, usePhotonCount=True)
        else:
            pts = np.array([[1.0, 2.0, 3.0], [1.0, 2.0 + 0.6 * 4.0, 3.0 + 0.8 * 4.0],
                            [1.0, 2.0 + 0.6 * 8.0, 3.0 + 0.8 * 8.0], [1.0, 2.0 + 0.6 * L, 3.0 + 0.8 * L]])
            t = 5.0 + np.array([0.0, 4.0, 8.0, L]) / 0.299792458
            track = light.ParticleTrack(4)
            track.setVertices(np.concatenate([pts, t[:, None]], 1))
            src = light.CherenkovTrackLightSource(track, usePhotonCount=True)
        rec = response.HitRecorder()
        rec.enableTags()
        rng = random.PhiloxRNG(key=0xCE4E)
        cb = trace.TrackRecordCallback(N, 4)
        tgt = target.InnerSphereTarget(position=(1.0, 5.0, 7.0), radius=60.0)
        tr = trace.VolumeForwardTracer(N, src, tgt, light.ConstWavelengthSource(lam), rec, rng, medium=store.media[med.name],
                                       nScattering=10, callback=cb, maxTime=float("inf"))
        cfg = {"tracer": tr, "response": rec, "callback": cb, "rng": rng, "store": store}
        H.runAny(cfg)
        total = float(H.hitView(cfg)["contrib"].astype(np.float64).sum()) / N
        tracks, lengths, _ = cb.result(0)
    ok = lengths >= 1
    d = tracks[ok, 1, :3].astype(np.float64) - tracks[ok, 0, :3]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    assert np.abs(d @ np.array([0.0, 0.6, 0.8]) - 1.0 / n).max() < 2e-4
    s = (tracks[ok, 0, :3].astype(np.float64) - np.array([1.0, 2.0, 3.0])) @ np.array([0.0, 0.6, 0.8])
    assert s.min() >= -1e-4 and s.max() <= L + 1e-4 and abs(s.mean() - L / 2) < 0.1
    assert np.allclose(tracks[ok, 0, 3], 5.0 + s / 0.299792458, atol=1e-3)  # emitted when the particle passes
    expected = 2.0 * np.pi * 7.2973525693 / (lam * 1e-3) ** 2 * (1.0 - 1.0 / n**2) * L
    print(f"{kind}: {total:.3f} vs {expected:.3f}")
    assert abs(total / expected - 1.0) < 0.01


def test_function_wavelength_source_samples_the_function():
    # sh/wavelengthsource.function.glsl: inverse CDF table; every sample carries the integral of fn
    from scipy import integrate, stats
    from theia_b200 import light
    fn = lambda lam: np.exp(-0.5 * ((lam - 480.0) / 60.0) ** 2) + 0.1  # noqa: E731
    lo, hi = 350.0, 650.0
    with _device.useDevice("cpu"):
        cfg = H.volumeScenario(1 << 14, lightKind="spherical", targetKind="inner", wavelength="function", resp="record")
        H.runOracle(cfg)
        lam = np.array(H.hitView(cfg)["wavelength"], dtype=np.float64)
        p = cfg["tracer"].wavelengthSource._firstBlock(0)
    assert p._contrib == pytest.approx(integrate.quad(fn, lo, hi)[0], rel=1e-5)
    assert lam.min() >= lo and lam.max() <= hi
    norm = integrate.quad(fn, lo, hi)[0]
    cdf = np.vectorize(lambda x: integrate.quad(fn, lo, x)[0] / norm)
    # one wavelength per path, many hits per path: test the distinct values
    assert stats.kstest(np.unique(lam), cdf).pvalue > 1e-3
