"""
The analytic acceptance tests the reference holds for the SCENE tracer (traversal -> processRayQuery -> Fresnel ->
MIS with a target guide -> response), ported with its scenes, parameters and tolerances:

  tests/test_trace_energy.py:58-175   test_SceneForwardTracer_GroundTruth   energy conservation, within 5 %
  tests/test_trace_energy.py:187-337  test_SceneForwardTracer_Crosscheck    target guide (MIS) against brute force, 2 % / 1 %
  tests/test_trace_energy.py:479-568  test_SceneForwardTracer_MultiMedia    nested media, reflection + transmission, 0.61 %
  tests/test_trace.py:1122-1193       test_volumeBorder                     no refraction at a volume border, group index
  tests/test_trace_energy.py:1341-1461 test_ScenePhotonTracer               photon counts against the forward tracer, 5 sigma
  tests/test_trace_energy.py:1464-1580 test_VolumePhotonTracer              the same for the volume tracers

Each runs on the CPU oracle (pinning the restatement; "not gpu" suite, reduced statistics with the tolerance widened
by the statistics only where the reference's bound is a statistical one) and on the CUDA path (-m gpu, at the
reference's statistics).
"""

import os

import numpy as np
import pytest

import harness as H
import theia_b200.units as u
from theia_b200 import _device, light, material, random, response, scene, target, trace

DEVICES = [pytest.param("cpu", id="oracle"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]
SPHERE = os.path.join(H.ASSETS, "sphere.stl")


def runBatches(tr, nBatches, collect):
    """runs nBatches batches of tracer tr (its rng advances on its own) and hands configuration i to collect"""
    cfg = {"tracer": tr, "response": tr.response, "callback": tr.callback, "rng": tr.rng}
    for b in range(nBatches):
        H.runAny(cfg, b % 2)
        collect(b % 2)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("mu_a,mu_s,g,polarized", [(0.0, 0.005, 0.0, False), (0.05, 0.01, 0.0, False),
                                                   (0.05, 0.01, -0.9, False), (0.05, 0.01, 0.9, False),
                                                   (0.0, 0.005, 0.0, True), (0.05, 0.01, -0.9, True)])
def test_SceneForwardTracer_GroundTruth(device, mu_a, mu_s, g, polarized):
    # sphere detector (mesh) filled with a scattering medium, spherical light source in its centre: no light escapes
    position, radius = (12.0, 15.0, 0.2), 100.0 * u.m
    budget, t0, lam = 1e9, 10.0 * u.ns, 400.0 * u.nm
    batch_size, n_batches = (2 * 1024 * 1024, 40) if device == "cuda" else (128 * 1024, 2)
    with _device.useDevice(device):
        model = H.EnergyModel(mu_a, mu_s, g)
        medium = model.createMedium()
        matStore = material.MaterialStore([material.Material("det", medium, None, flags="DB")])
        meshStore = scene.MeshStore({"sphere": SPHERE})
        det = meshStore.createInstance("sphere", "det", scene.Transform.TRS(scale=radius, translate=position), detectorId=0)
        sc = scene.Scene([det], matStore.material, medium=matStore.media[medium.name])
        photons = light.UniformWavelengthSource(lambdaRange=(lam, lam))
        src = light.SphericalLightSource(position=position, timeRange=(t0, t0), budget=budget)
        rng = random.PhiloxRNG(key=0xC0FFEE)
        recorder = response.HitRecorder(polarized=polarized)
        tracer = trace.SceneForwardTracer(batch_size, src, photons, recorder, rng, scene=sc, maxPathLength=10,
                                          scatterCoefficient=0.05, maxTime=float("inf"), polarized=polarized)
        rng.autoAdvance = tracer.nRNGSamples
        total, posSum, nHits, variances, stokes = [0.0], [np.zeros(3)], [0], [], []
        vg = 1.0 / model.ng * u.c

        def collect(i):
            v = recorder.queue.view(i)
            n = min(v.count, recorder.capacity)
            time, value = v["time"][:n].astype(np.float64), v["contrib"][:n].astype(np.float64)
            total[0] += float((value * np.exp(mu_a * vg * (time - t0))).sum())  # undo the attenuation
            pos = np.array(v["position"][:n], dtype=np.float64)
            posSum[0] += pos.sum(0)
            nHits[0] += n
            variances.append(pos.var(0))
            if polarized:
                stokes.append(np.array(v["stokes"][:n]))

        runBatches(tracer, n_batches, collect)
    estimate = total[0] / (batch_size * n_batches)
    print(f"estimate / budget - 1 = {estimate / budget - 1.0:+.4f}")
    assert np.abs(estimate / budget - 1.0) < 0.05
    # "estimate < budget" (biased by ignoring longer paths) is a statement about the expectation: the importance sampled
    # free paths give single hits up to 1e4 times the budget, so it only shows at the reference's 8e7 paths
    if device == "cuda":
        assert estimate < budget
    # hits uniform on the (unit, object space) sphere
    assert np.abs(posSum[0] / nHits[0]).max() < (5e-3 if device == "cuda" else 1.5e-2)
    assert np.abs(np.vstack(variances) - 1 / 3).max() < (0.01 if device == "cuda" else 0.02)
    if polarized:
        s = np.concatenate(stokes)
        assert np.abs(s[:, 0] - 1.0).max() < 1e-5
        assert s[:, 1:].max() <= 1.0 and s[:, 1:].min() >= -1.0
        assert np.all(np.square(s[:, 1:]).sum(-1) <= 1.0 + 1e-6)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("mu_a,mu_s,g,polarized", [(0.0, 0.01, 0.0, False), (0.05, 0.005, 0.5, False),
                                                   (0.0, 0.01, 0.0, True), (0.05, 0.005, -0.5, True)])
def test_SceneForwardTracer_Crosscheck(device, mu_a, mu_s, g, polarized):
    # target sampling (MIS with a SphereTargetGuide) against the tracer without it
    position, radius, light_pos = (0.0, 0.0, 0.0), 5.0 * u.m, (-6.0, 0.0, 0.0)
    budget, t0, lam = 1e9, 30.0 * u.ns, 400.0 * u.nm
    max_length, scatter_coef, maxTime = 10, 0.01, 600.0
    batch_size, n_batches = (1024 * 1024, 100) if device == "cuda" else (256 * 1024, 12)
    bin_t0, bin_size, n_bins = 0.0, 20.0, 30
    with _device.useDevice(device):
        model = H.EnergyModel(mu_a, mu_s, g)
        medium = model.createMedium()
        matStore = material.MaterialStore([material.Material("det", None, medium, flags="DB")])
        meshStore = scene.MeshStore({"sphere": SPHERE})
        guide = target.SphereTargetGuide(position=position, radius=radius)
        det = meshStore.createInstance("sphere", "det", scene.Transform.TRS(scale=radius, translate=position), detectorId=0)
        sc = scene.Scene([det], matStore.material, medium=matStore.media[medium.name])
        photons = light.UniformWavelengthSource(lambdaRange=(lam, lam))
        src = light.SphericalLightSource(position=light_pos, timeRange=(t0, t0), budget=budget)
        rng = random.PhiloxRNG(key=0xC0FFEE)

        def lightCurve(pathLength, targetGuide):
            resp = response.HistogramHitResponse(response.UniformValueResponse(), nBins=n_bins, binSize=bin_size, t0=bin_t0,
                                                 normalization=1 / batch_size)
            tracer = trace.SceneForwardTracer(batch_size, src, photons, resp, rng, scene=sc, maxPathLength=pathLength,
                                              targetGuide=targetGuide, scatterCoefficient=scatter_coef, maxTime=maxTime,
                                              polarized=polarized)
            rng.autoAdvance = tracer.nRNGSamples
            hists = []
            runBatches(tracer, n_batches, lambda i: hists.append(resp.result(i).astype(np.float64).copy()))
            return np.mean(hists, 0)

        truth_hist = lightCurve(2 * max_length, None)  # higher variance -> more samples
        hist = lightCurve(max_length, guide)
    truth, estimate = truth_hist.sum(), hist.sum()
    print(f"estimate / truth - 1 = {estimate / truth - 1.0:+.4f}")
    assert truth > 0.0
    # the reference's 2 % holds at its 1e8 paths; the brute force side is the noisy one (hits of a 5 m sphere from
    # scattered light): 40 times fewer paths on the oracle widen it
    assert abs(estimate / truth - 1.0) < (0.02 if device == "cuda" else 0.08)
    with np.errstate(divide="ignore", invalid="ignore"):
        log_err = (np.log10(hist) - np.log10(truth_hist)) / np.log10(hist)
    log_err = np.nan_to_num(log_err, nan=0.0, posinf=0.0, neginf=0.0)
    print(f"mean log error {np.abs(log_err).mean():.4f}")
    assert np.abs(log_err).mean() < (0.01 if device == "cuda" else 0.04)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("polarized", [True, False])
def test_SceneForwardTracer_MultiMedia(device, polarized):
    # spherical light source inside a spherical detector with an inner and an outer shell of different media: every
    # crossing of the inner sphere reflects or transmits (importance sampled), all energy must arrive
    position, light_pos = (12.0, 15.0, 0.2), (2.0, 5.0, 0.0)
    radius, radius_inner = 100.0 * u.m, 65.0 * u.m
    budget, t0, lam = 1e9, 10.0 * u.ns, 400.0 * u.nm
    batch_size, n_batches = (2 * 1024 * 1024, 10) if device == "cuda" else (128 * 1024, 2)
    with _device.useDevice(device):
        media_inner = H.EnergyModel(0.0, 0.01, 0.9, n=1.8).createMedium(name="inner")
        media_outer = H.EnergyModel(0.0, 0.04, 0.9, n=1.2).createMedium(name="outer")
        mat_det = material.Material("det", media_outer, None, flags="DB")
        mat_inner = material.Material("inner", media_inner, media_outer, flags="TR")
        matStore = material.MaterialStore([mat_det, mat_inner])
        meshStore = scene.MeshStore({"sphere": SPHERE})
        inner = meshStore.createInstance("sphere", "inner", scene.Transform.TRS(scale=radius_inner, translate=position))
        det = meshStore.createInstance("sphere", "det", scene.Transform.TRS(scale=radius, translate=position))
        sc = scene.Scene([inner, det], matStore.material)
        photons = light.ConstWavelengthSource(lam)
        src = light.SphericalLightSource(position=light_pos, timeRange=(t0, t0), budget=budget)
        rng = random.PhiloxRNG(key=0xC0FFEE)
        resp = response.HistogramHitResponse(response.UniformValueResponse(), nBins=400, binSize=100.0 * u.ns)
        tracer = trace.SceneForwardTracer(batch_size, src, photons, resp, rng, sc, maxPathLength=50, scatterCoefficient=0.03,
                                          sourceMedium=matStore.media["inner"], maxTime=float("inf"), polarized=polarized)
        rng.autoAdvance = tracer.nRNGSamples
        hists = []
        runBatches(tracer, n_batches, lambda i: hists.append(resp.result(i).astype(np.float64).copy()))
    estimate = np.mean(hists, 0).sum()
    print(f"estimate / budget - 1 = {estimate / budget - 1.0:+.5f}")
    assert abs(estimate / budget - 1.0) < 0.0061


@pytest.mark.parametrize("device", DEVICES)
def test_volumeBorder(device):
    # a laser from vacuum into glass marked as volume border: no refraction, but the ray slows down
    N, LAMBDA = 32 * 256, 500.0 * u.nm
    with _device.useDevice(device):
        model = material.BK7Model()
        glass = model.createMedium()
        matStore = material.MaterialStore([material.Material("mat", glass, None, flags=("V", "B"))])
        store = scene.MeshStore({"cube": os.path.join(H.ASSETS, "cube.ply")})
        cube = store.createInstance("cube", "mat", scene.Transform.TRS(scale=50.0, translate=(75.0, 0.0, 0.0)))
        sc = scene.Scene([cube], matStore.material)
        rng = random.PhiloxRNG(key=0xC01DC0FFEE)
        photons = light.UniformWavelengthSource(lambdaRange=(LAMBDA, LAMBDA))
        # hits the cube at an angle: a refraction would show
        source = light.PencilLightSource(position=(0.0, 0.0, 0.0), direction=(0.8, 0.36, 0.48), timeRange=(0.0, 0.0), budget=1.0)
        tracker = trace.TrackRecordCallback(N, 4)
        tracer = trace.SceneForwardTracer(N, source, photons, response.EmptyResponse(), rng, callback=tracker, scene=sc,
                                          targetId=1, maxPathLength=2)
        H.runAny({"tracer": tracer, "response": tracer.response, "callback": tracker, "rng": rng})
        track, lengths, codes = tracker.result(0)
    assert lengths.min() >= 2
    track = track.astype(np.float64)
    d1 = track[:, 1, :3] - track[:, 0, :3]
    l1 = np.sqrt(np.square(d1).sum(-1))
    d2 = track[:, 2, :3] - track[:, 1, :3]
    l2 = np.sqrt(np.square(d2).sum(-1))
    cos_theta = np.multiply(d1, d2).sum(-1) / (l1 * l2)
    assert cos_theta.min() >= 1.0 - 1e-5
    t1 = track[:, 1, 3] - track[:, 0, 3]
    t2 = track[:, 2, 3] - track[:, 1, 3]
    n1 = material.speed_of_light / (l1 / t1)
    n2 = material.speed_of_light / (l2 / t2)
    n2_exp = material.speed_of_light / model.group_velocity(LAMBDA)
    assert np.abs(n1 - 1.0).max() < 5e-5
    assert np.abs(n2 - n2_exp).max() < 5e-5


PHOTON_CASES = [(0.0, 0.05, 0.0, False), (0.05, 0.05, 0.0, False), (0.05, 0.01, -0.9, False), (0.05, 0.01, 0.9, False),
                (0.05, 0.01, -0.9, True)]  # (the reference leaves (0.0, 0.05, 0.0, True) out: "polarization seems broken")


def _photonAgainstForward(device, makeForward, makePhoton, mu_a, budget, batch_size, n_batches):
    """The forward tracer's light curve, summed, is the number of photons the photon tracer must detect (5 sigma)."""
    hist_tracer, hist_response, rng = makeForward()
    rng.autoAdvance = hist_tracer.nRNGSamples
    hists = []
    runBatches(hist_tracer, n_batches, lambda i: hists.append(np.array(hist_response.result(i), dtype=np.float64)))
    per_batch = np.array([h.sum() for h in hists])
    exp_hits = per_batch.mean()
    tracer, resp = makePhoton(rng)
    H.runAny({"tracer": tracer, "response": resp, "callback": tracer.callback, "rng": rng})
    hits = resp.result(0)
    assert hits is not None
    print(f"photon tracer {hits.count} hits, forward tracer expects {exp_hits:.1f}")
    if mu_a != 0.0:
        # the reference's bound; at the reduced statistics of the oracle run the error of the forward estimate itself
        # (standard error over its batches) is not negligible next to it
        slack = 0.0 if device == "cuda" else 4.0 * per_batch.std(ddof=1) / np.sqrt(len(per_batch))
        assert np.abs(hits.count - exp_hits) < 5.0 * np.sqrt(exp_hits) + slack
    else:
        assert hits.count == int(budget)  # no absorption: every photon is recovered


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("mu_a,mu_s,g,polarized", PHOTON_CASES)
def test_ScenePhotonTracer(device, mu_a, mu_s, g, polarized):
    position, radius = (12.0, 15.0, 0.2), 60.0 * u.m
    t0, lam = 10.0 * u.ns, 400.0 * u.nm
    budget, batch_size, n_batches = (1e6, 1024 * 1024, 20) if device == "cuda" else (1e5, 32 * 1024, 8)
    with _device.useDevice(device):
        model = H.EnergyModel(mu_a, mu_s, g)
        medium = model.createMedium()
        matStore = material.MaterialStore([material.Material("det", medium, None, flags="DB")])
        meshStore = scene.MeshStore({"sphere": SPHERE})
        det = meshStore.createInstance("sphere", "det", scene.Transform.TRS(scale=radius, translate=position), detectorId=0)
        sc = scene.Scene([det], matStore.material, medium=matStore.media[medium.name])
        photons = light.UniformWavelengthSource(lambdaRange=(lam, lam))
        src = light.SphericalLightSource(position=position, timeRange=(t0, t0), budget=budget)

        def makeForward():
            rng = random.PhiloxRNG(key=0xC01DC0FFEE)
            hist = response.HistogramHitResponse(response.UniformValueResponse(), binSize=50.0 * u.ns, nBins=400)
            tr = trace.SceneForwardTracer(batch_size, src, photons, hist, rng, sc, maxPathLength=80, scatterCoefficient=0.05,
                                          maxTime=float("inf"), polarized=polarized)
            return tr, hist, rng

        def makePhoton(rng):
            resp = response.StoreTimeHitResponse(response.UniformValueResponse())
            tr = trace.ScenePhotonTracer(int(budget), src, photons, resp, rng, sc, nScatteringPerRun=5, nRuns=20,
                                         maxTime=float("inf"), polarized=polarized, callback=trace.EventStatisticCallback())
            return tr, resp

        _photonAgainstForward(device, makeForward, makePhoton, mu_a, budget, batch_size, n_batches)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("mu_a,mu_s,g,polarized", PHOTON_CASES)
def test_VolumePhotonTracer(device, mu_a, mu_s, g, polarized):
    position, radius = (12.0, 15.0, 0.2), 60.0 * u.m
    t0, lam = 10.0 * u.ns, 400.0 * u.nm
    budget, batch_size, n_batches = (1e6, 1024 * 1024, 20) if device == "cuda" else (1e5, 32 * 1024, 8)
    with _device.useDevice(device):
        model = H.EnergyModel(mu_a, mu_s, g)
        medium = model.createMedium()
        store = material.MaterialStore([], media=[medium])
        photons = light.UniformWavelengthSource(lambdaRange=(lam, lam))
        src = light.SphericalLightSource(position=position, timeRange=(t0, t0), budget=budget)
        tgt = target.InnerSphereTarget(position=position, radius=radius)

        def makeForward():
            rng = random.PhiloxRNG(key=0xC0FFEE)
            hist = response.HistogramHitResponse(response.UniformValueResponse(), binSize=50.0 * u.ns, nBins=400)
            tr = trace.VolumeForwardTracer(batch_size, src, tgt, photons, hist, rng, medium=store.media[medium.name],
                                           maxTime=float("inf"), nScattering=80, scatterCoefficient=0.05, polarized=polarized)
            return tr, hist, rng

        def makePhoton(rng):
            resp = response.StoreTimeHitResponse(response.UniformValueResponse())
            tr = trace.VolumePhotonTracer(int(budget), src, tgt, photons, resp, rng, medium=store.media[medium.name],
                                          nScatteringPerRun=5, nRuns=40, maxTime=float("inf"), polarized=polarized,
                                          callback=trace.EventStatisticCallback())
            return tr, resp

        _photonAgainstForward(device, makeForward, makePhoton, mu_a, budget, batch_size, n_batches)
