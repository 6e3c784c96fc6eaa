"""
The flag sweeps of the reference's tracer tests, ported with their scenes and assertions:
  tests/test_trace.py:82-212   test_VolumeForwardTracer       2^5 combinations
  tests/test_trace.py:335-457  test_SceneForwardTracer        2^7 combinations
  tests/test_trace.py:460-518  test_SceneForwardTracer_multiTarget
  tests/test_trace.py:947-1046 test_EventStatisticCallback    every result code of a scene with all material kinds
  tests/test_trace.py:1048-1119 test_TrackRecordCallback
  tests/test_trace.py:1466-1561 test_ScenePhotonTracer, :1563-1654 test_VolumePhotonTracer
Every combination runs on the CPU oracle ("not gpu" suite) and on the CUDA path (-m gpu).
"""

import os

import numpy as np
import pytest

import harness as H
import theia_b200.units as u
from theia_b200 import _device, light, material, random, response, scene, target, trace
from theia_b200.testing import WaterTestModel

DEVICES = [pytest.param("cpu", id="oracle"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]
BOOL = [True, False]
R_SCALE = 0.99547149974733  # radius of the sphere inscribed in assets/sphere.stl (icosphere)


def run(tracer):
    H.runAny({"tracer": tracer, "response": tracer.response, "callback": tracer.callback, "rng": tracer.rng})


def checkPolarization(hits):
    assert np.abs(hits["stokes"][:, 0] - 1.0).max() < 1e-6
    assert np.sqrt(np.square(hits["stokes"][:, 1:]).sum(-1).max()) - 1.0 <= 1e-6
    assert hits["stokes"][:, 1:].max() <= 1.0 and hits["stokes"][:, 1:].min() >= -1.0
    # the reference frame is perpendicular to the plane of incidence and normalised
    polRef = hits["polarizationRef"]
    assert np.abs(np.square(polRef).sum(-1) - 1.0).max() < 1e-6
    polRef_exp = np.cross(hits["direction"], hits["normal"])
    d = np.square(polRef_exp).sum(-1)
    mask = d > 1e-7  # for very small angles the old frame is kept
    polRef_exp = polRef_exp[mask] / np.sqrt(d[mask])[:, None]
    assert np.abs(np.abs((polRef_exp * polRef[mask]).sum(-1)) - 1.0).max() < 1e-6


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("disableDirect", BOOL)
@pytest.mark.parametrize("disableTarget", BOOL)
@pytest.mark.parametrize("disableScattering", BOOL)
@pytest.mark.parametrize("limitTime", BOOL)
@pytest.mark.parametrize("polarized", BOOL)
def test_VolumeForwardTracer(device, disableDirect, disableTarget, disableScattering, limitTime, polarized):
    N, N_SCATTER = 32 * 256, 6
    T0, T1 = 10.0 * u.ns, 20.0 * u.ns
    T_MAX = 1.0 * u.us if limitTime else 100.0 * u.us
    capacity, objectId = 48 * 256, 42
    light_pos, light_budget = (-1.0, -7.0, 0.0), 1000.0
    target_pos, target_radius = (5.0, 2.0, -8.0), 4.0 * u.m
    with _device.useDevice(device):
        water = WaterTestModel().createMedium()
        store = material.MaterialStore([], media=[water])
        d_min = np.sqrt(np.square(np.subtract(light_pos, target_pos)).sum(-1)) - target_radius
        t_min = d_min / np.max(water.group_velocity) + T0
        rng = random.PhiloxRNG(key=0xC01DC0FFEE)
        photons = light.UniformWavelengthSource()
        source = light.SphericalLightSource(position=light_pos, budget=light_budget, timeRange=(T0, T1))
        tgt = target.SphereTarget(position=target_pos, radius=target_radius)
        recorder = response.HitRecorder(polarized=polarized)
        stats = trace.EventStatisticCallback()
        tracer = trace.VolumeForwardTracer(N, source, tgt, photons, recorder, rng, medium=store.media["water"], objectId=objectId,
                                           capacity=capacity, nScattering=N_SCATTER,
                                           scatterCoefficient=0.0 if disableScattering else float("NaN"), callback=stats,
                                           maxTime=T_MAX, polarized=polarized, disableDirectLighting=disableDirect,
                                           disableTargetSampling=disableTarget)
        run(tracer)
        hits = recorder.queue.view(0)
        if disableDirect and disableScattering:
            assert hits.count == 0 and stats.scattered == 0  # this combination cannot create hits
            return
        assert 0 < hits.count <= tracer.maxHits
        hits = hits[: hits.count]
        assert np.allclose(np.square(hits["position"]).sum(-1), 1.0)
        assert np.allclose(np.square(hits["normal"]).sum(-1), 1.0)
        assert np.min(hits["time"]) >= t_min and np.max(hits["time"]) <= T_MAX
        assert np.allclose(hits["objectId"], objectId)
        if polarized:
            checkPolarization(hits)
        assert stats.created == tracer.batchSize
        if disableScattering:
            # The reference asserts scattered == 0. A detected path goes on (RESULT_CODE_RAY_DETECTED > 0) from ON the sphere,
            # where intersectTarget may find the surface again at t ~ 1e-7 with prob = inf -> valid = false, so the free path
            # is clamped to that t and the step counts as a scatter event (sh/target.sphere.glsl:45-78 +
            # sh/tracer.volume.forward.glsl:163-167; whether t rounds to +- is a coin toss in any float32 implementation).
            # Only such re-intersections may show up here.
            assert stats.scattered <= stats.detected
        else:
            assert stats.scattered > 0
        if disableTarget:
            assert stats.detected > 0
            assert (stats.absorbed > 0) if disableDirect else (stats.absorbed == 0)
        else:
            assert stats.absorbed > 0 and stats.detected == 0


def sweepScene(device):
    """two glass filled detector spheres (ids 0, 1) in water; light source next to them (tests/test_trace.py:359-385)"""
    water = WaterTestModel().createMedium()
    glass = material.BK7Model().createMedium()
    matStore = material.MaterialStore([material.Material("mat", glass, water, flags=("DR", "B"))])
    store = scene.MeshStore({"cube": os.path.join(H.ASSETS, "cone.stl"), "sphere": os.path.join(H.ASSETS, "sphere.stl")})
    r, d = 40.0 * u.m, 5.0 * u.m
    x, y, z = 10.0, 5.0, -5.0
    c1 = store.createInstance("sphere", "mat", scene.Transform.TRS(scale=r, translate=(x, y, z + r + d)), detectorId=0)
    c2 = store.createInstance("sphere", "mat", scene.Transform.TRS(scale=r, translate=(x, y, z - r - d), rotate=(0.0, 0.0, 1.0, 110.0)),
                              detectorId=1)
    sc = scene.Scene([c1, c2], matStore.material, medium=matStore.media["water"])
    sc._keepAlive = (matStore, store)  # the scene holds addresses only (as in the reference): the stores must outlive it
    return water, sc, (x, y, z - r - d), r


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("useRefractedHitDir", BOOL)
@pytest.mark.parametrize("disableDirect", BOOL)
@pytest.mark.parametrize("disableVolumeBorder", BOOL)
@pytest.mark.parametrize("disableTransmission", BOOL)
@pytest.mark.parametrize("disableTarget", BOOL)
@pytest.mark.parametrize("disableScattering", BOOL)
@pytest.mark.parametrize("polarized", BOOL)
def test_SceneForwardTracer(device, polarized, disableDirect, disableVolumeBorder, disableTransmission, disableTarget,
                            disableScattering, useRefractedHitDir):
    N, MAX_PATH = 32 * 256, 10
    T0, T1, T_MAX = 10.0 * u.ns, 20.0 * u.ns, 1.0 * u.us
    light_pos = (-1.0, -7.0, 0.0)
    with _device.useDevice(device):
        water, sc, target_pos, r = sweepScene(device)
        guide = target.SphereTargetGuide(position=target_pos, radius=r)
        d_min = np.sqrt(np.square(np.subtract(target_pos, light_pos)).sum(-1)) - r
        t_min = d_min / np.max(water.group_velocity) + T0
        rng = random.PhiloxRNG(key=0xC01DC0FFEE)
        source = light.SphericalLightSource(position=light_pos, timeRange=(T0, T1), budget=1000.0)
        recorder = response.HitRecorder(polarized=polarized)
        stats = trace.EventStatisticCallback()
        tracer = trace.SceneForwardTracer(N, source, light.UniformWavelengthSource(), recorder, rng, sc, capacity=48 * 256,
                                          maxPathLength=MAX_PATH, targetId=1, targetGuide=None if disableTarget else guide,
                                          maxTime=T_MAX, polarized=polarized, callback=stats,
                                          scatterCoefficient=0.0 if disableScattering else float("NaN"),
                                          disableDirectLighting=disableDirect, disableTransmission=disableTransmission,
                                          disableVolumeBorder=disableVolumeBorder, useRefractedHitDir=useRefractedHitDir)
        run(tracer)
        hits = recorder.queue.view(0)
        assert 0 < hits.count <= tracer.maxHits
        hits = hits[: hits.count]
        assert stats.created == N
        if disableScattering:
            assert stats.scattered == 0
        r_hits = np.sqrt(np.square(hits["position"]).sum(-1))
        assert np.all((r_hits <= 1.0 + 1e-6) & (r_hits >= R_SCALE - 1e-6))
        assert np.allclose(np.square(hits["normal"]).sum(-1), 1.0)
        assert np.min(hits["time"]) >= t_min and np.max(hits["time"]) <= T_MAX
        assert np.all(hits["objectId"] == 1)
        if polarized:
            checkPolarization(hits)


@pytest.mark.parametrize("device", DEVICES)
def test_SceneForwardTracer_multiTarget(device):
    N, T0, T1, T_MAX = 32 * 256, 10.0 * u.ns, 20.0 * u.ns, 1.0 * u.us
    with _device.useDevice(device):
        water, sc, _, _ = sweepScene(device)
        rng = random.PhiloxRNG(key=0xC01DC0FFEE)
        source = light.SphericalLightSource(position=(-1.0, -7.0, 0.0), timeRange=(T0, T1), budget=1000.0)
        recorder = response.HitRecorder()
        tracer = trace.SceneForwardTracer(N, source, light.UniformWavelengthSource(), recorder, rng, sc, capacity=48 * 256,
                                          targetId=-1, maxPathLength=10, maxTime=T_MAX, callback=trace.EventStatisticCallback())
        run(tracer)
        hits = recorder.queue.view(0)
        hits = hits[: hits.count]
    r_hits = np.sqrt(np.square(hits["position"]).sum(-1))
    assert np.all((r_hits <= 1.0 + 1e-6) & (r_hits >= R_SCALE - 1e-6))
    assert np.allclose(np.square(hits["normal"]).sum(-1), 1.0)
    assert (hits["objectId"] == 0).sum() > 0 and (hits["objectId"] == 1).sum() > 0


@pytest.mark.parametrize("device", DEVICES)
def test_EventStatisticCallback(device):
    N, T0, T1 = 32 * 1024, 10.0 * u.ns, 20.0 * u.ns
    with _device.useDevice(device):
        water = WaterTestModel().createMedium()
        glass = material.BK7Model().createMedium()
        matStore = material.MaterialStore([material.Material("mat", glass, water, flags=("DR", "B")),
                                           material.Material("vol", glass, water, flags="V"),
                                           material.Material("abs", glass, water, flags="B")])
        store = scene.MeshStore({"sphere": os.path.join(H.ASSETS, "sphere.stl")})
        r = 20.0
        inst = [store.createInstance("sphere", "mat", scene.Transform.TRS(scale=r, translate=(17.0, 17.0, 17.0)), detectorId=0),
                store.createInstance("sphere", "mat", scene.Transform.TRS(scale=r, translate=(-17.0, -17.0, -17.0)), detectorId=1),
                store.createInstance("sphere", "vol", scene.Transform.TRS(scale=r, translate=(17.0, 17.0, -17.0))),
                store.createInstance("sphere", "abs", scene.Transform.TRS(scale=r, translate=(-17.0, 17.0, 17.0)))]
        bbox = scene.RectBBox((-50.0,) * 3, (50.0,) * 3)
        sc = scene.Scene(inst, matStore.material, medium=matStore.media["water"], bbox=bbox)
        sc._keepAlive = (matStore, store)
        rng = random.PhiloxRNG(key=0xC01DC0FFEE)
        stats = trace.EventStatisticCallback()
        tracer = trace.SceneForwardTracer(N, light.SphericalLightSource(timeRange=(T0, T1)), light.UniformWavelengthSource(),
                                          response.EmptyResponse(), rng, callback=stats, scene=sc, targetId=1, maxPathLength=1,
                                          scatterCoefficient=0.005 / u.m, maxTime=100.0 * u.ns)
        names = ("created", "absorbed", "hit", "detected", "scattered", "lost", "decayed", "volume", "maxIter", "error", "mismatch")
        assert all(getattr(stats, n) == 0 for n in names)  # starts empty
        run(tracer)
        assert stats.created == N
        for n in ("absorbed", "hit", "detected", "scattered", "lost", "decayed", "volume", "maxIter"):
            assert getattr(stats, n) > 0, n
        assert stats.error == 0 and stats.mismatch == 0
        total = stats.absorbed + stats.hit + stats.detected + stats.scattered + stats.lost + stats.decayed + stats.volume
        assert total == N
        stats.reset()
        assert all(getattr(stats, n) == 0 for n in names)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("polarized", BOOL)
@pytest.mark.parametrize("polarizedTrack", BOOL)
def test_TrackRecordCallback(device, polarizedTrack, polarized):
    N, N_SCATTER = 32 * 256, 6
    LENGTH = N_SCATTER + 2  # one more than needed
    T0, T1, T_MAX = 10.0 * u.ns, 20.0 * u.ns, 500.0 * u.ns
    target_pos, target_radius = (5.0, 2.0, -8.0), 4.0 * u.m
    light_pos, light_dir = (100.0, -50.0, 20.0), (1.0, 0.0, 0.0)
    stokes, polRef = (1.0, 0.5, 0.1, 0.4), (0.0, 1.0, 0.0)
    with _device.useDevice(device):
        water = WaterTestModel().createMedium()
        store = material.MaterialStore([], media=[water])
        rng = random.PhiloxRNG(key=0xC01DC0FFEE)
        source = light.PencilLightSource(position=light_pos, direction=light_dir, timeRange=(T0, T1), stokes=stokes,
                                         polarizationRef=polRef)
        track = trace.TrackRecordCallback(N, LENGTH, polarized=polarizedTrack)
        tracer = trace.VolumeForwardTracer(N, source, target.SphereTarget(position=target_pos, radius=target_radius),
                                           light.UniformWavelengthSource(), response.EmptyResponse(), rng, callback=track,
                                           medium=store.media["water"], nScattering=N_SCATTER, scatterCoefficient=0.1,
                                           maxTime=T_MAX, traceBBox=scene.RectBBox((-200.0 * u.m,) * 3, (200.0 * u.m,) * 3),
                                           polarized=polarized)
        run(tracer)
        tracks, lengths, codes = track.result(0)
    assert tracks.shape == (N, LENGTH, 11 if polarizedTrack else 4)
    assert np.all(tracks[:, 0, :3] == light_pos)
    assert tracks[:, 0, 3].min() >= T0 and tracks[:, 0, 3].max() <= T1
    # pencil beam in x direction: y and z of the first scatter are still deterministic
    assert np.all(tracks[:, 1, 1] == light_pos[1]) and np.all(tracks[:, 1, 2] == light_pos[2])
    assert lengths.min() >= 1 and lengths.max() <= N_SCATTER
    assert codes.min() >= min(c.value for c in trace.EventResultCode)
    assert codes.max() <= max(c.value for c in trace.EventResultCode)
    if polarizedTrack and polarized:
        assert np.isclose(tracks[:, 0, 4:8], stokes).all()  # only the first state is known
        assert np.isclose(tracks[:, 0, 8:], polRef).all()
    elif polarizedTrack:
        assert np.isclose(tracks[:, 0, 4:], (1.0, *(0.0,) * 6)).all()  # unpolarised with zero reference


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("polarized", BOOL)
def test_ScenePhotonTracer(device, polarized):
    N, N_RUNS, N_SCATTER = 32 * 256, 5, 5
    T0, T1, T_MAX = 10.0 * u.ns, 20.0 * u.ns, 1.0 * u.ms
    with _device.useDevice(device):
        water, sc, _, _ = sweepScene(device)
        rng = random.PhiloxRNG(key=0xC01DC0FFEE)
        source = light.SphericalLightSource(position=(-1.0, -7.0, 0.0), timeRange=(T0, T1), budget=float(N))
        recorder = response.HitRecorder(polarized=polarized)
        stats = trace.EventStatisticCallback()
        tracer = trace.ScenePhotonTracer(N, source, light.UniformWavelengthSource(), recorder, rng, sc, capacity=48 * 256,
                                         targetId=-1, nRuns=N_RUNS, nScatteringPerRun=N_SCATTER, maxTime=T_MAX, callback=stats,
                                         polarized=polarized)
        run(tracer)
        hits = recorder.queue.view(0)
        assert 0 < hits.count <= tracer.maxHits
        hits = hits[: hits.count]
        assert stats.created == N
    r_hits = np.sqrt(np.square(hits["position"]).sum(-1))
    assert np.all((r_hits <= 1.0 + 1e-6) & (r_hits >= R_SCALE - 1e-6))
    assert np.allclose(np.square(hits["normal"]).sum(-1), 1.0)
    assert np.max(hits["time"]) <= T_MAX
    assert (hits["objectId"] == 0).sum() > 0 and (hits["objectId"] == 1).sum() > 0
    if polarized:
        checkPolarization(hits)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("polarized", BOOL)
def test_VolumePhotonTracer(device, polarized):
    N, N_RUNS, N_SCATTER = 32 * 256, 5, 5
    T0, T1 = 10.0 * u.ns, 20.0 * u.ns
    objectId, light_pos = 42, (-1.0, -7.0, 0.0)
    target_pos, target_radius = (5.0, 2.0, -8.0), 4.0 * u.m
    with _device.useDevice(device):
        water = WaterTestModel().createMedium()
        store = material.MaterialStore([], media=[water])
        d_min = np.sqrt(np.square(np.subtract(light_pos, target_pos)).sum(-1)) - target_radius
        t_min = d_min / np.max(water.group_velocity) + T0
        rng = random.PhiloxRNG(key=0xC01DC0FFEE)
        source = light.SphericalLightSource(position=light_pos, budget=1000.0, timeRange=(T0, T1))
        recorder = response.HitRecorder(polarized=polarized)
        stats = trace.EventStatisticCallback()
        tracer = trace.VolumePhotonTracer(N, source, target.SphereTarget(position=target_pos, radius=target_radius),
                                          light.UniformWavelengthSource(), recorder, rng, medium=store.media["water"],
                                          objectId=objectId, capacity=48 * 256, nRuns=N_RUNS, nScatteringPerRun=N_SCATTER,
                                          maxTime=float("inf"), callback=stats, polarized=polarized)
        run(tracer)
        hits = recorder.queue.view(0)
        assert 0 < hits.count <= N
        hits = hits[: hits.count]
        assert stats.created == N
    assert np.allclose(np.square(hits["position"]).sum(-1), 1.0)
    assert np.allclose(np.square(hits["normal"]).sum(-1), 1.0)
    assert np.min(hits["time"]) >= t_min
    assert np.allclose(hits["objectId"], objectId)
    if polarized:
        checkPolarization(hits)
