"""
Shared scenario builders and the parity harness of the test-suite.

Every scenario is built twice from the same Python description: once on the "cuda" device
(the product: CUDA kernels behind the C-ABI) and once under ``useDevice("cpu")`` for the CPU
oracle (``oracle/``), which consumes the same launch description over host memory. The
scenarios are the configurations of BASELINE.json / SURVEY.md section 8d at test sizes.
"""

from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import theia_b200.units as u  # noqa: E402
from theia_b200 import _device, light, material, random, response, scene, target, trace  # noqa: E402
from theia_b200.testing import WaterTestModel  # noqa: E402

ASSETS = os.path.join(ROOT, "assets")


class HomogenousModel(material.DispersionFreeMedium, material.HenyeyGreensteinPhaseFunction, material.MediumModel):
    """examples/01_volume_tracing.ipynb cell 5 of the reference"""

    ModelName = "homogenous"

    def __init__(self, *, n=1.33, ng=1.33, mu_a=0.05, mu_s=0.02, g=0.2) -> None:
        material.DispersionFreeMedium.__init__(self, n=n, ng=ng, mu_a=mu_a, mu_s=mu_s)
        material.HenyeyGreensteinPhaseFunction.__init__(self, g)


def makeResponse(kind: str, *, polarized: bool = False, nBins: int = 100, binSize: float = 5.0, t0: float = 0.0,
                 capacity: int = 0, **kw):
    if kind == "record":
        r = response.HitRecorder(polarized=polarized, capacity=capacity)
        r.enableTags()
        return r
    if kind == "hist":
        return response.HistogramHitResponse(response.UniformValueResponse(), nBins=nBins, binSize=binSize, t0=t0)
    if kind == "empty":
        return response.EmptyResponse()
    if kind == "kernel":
        return response.KernelHistogramHitResponse(response.UniformValueResponse(), nBins=nBins, binSize=binSize, t0=t0,
                                                   kernelBandwidth=0.8 * binSize, kernelSupport=2.5 * binSize)
    if kind == "time":
        r = response.StoreTimeHitResponse(response.UniformValueResponse(), capacity=capacity)
        r.enableTags()
        return r
    if kind == "objhist":
        return response.ObjectHistogramHitResponse(response.UniformValueResponse(), nObjects=kw["nObjects"], nBins=nBins,
                                                   binSize=binSize, t0=t0)
    raise ValueError(kind)


def makeCallback(kind: str | None, capacity: int, length: int, polarized: bool = False):
    if kind == "stat":
        return trace.EventStatisticCallback()
    if kind == "track":
        return trace.TrackRecordCallback(capacity, length, polarized=polarized)
    return None


def hostSourceData(n: int, *, polarized: bool, seed: int = 11, origin=(1.0, 0.0, 0.0), axis=(-1.0, 0.0, 0.0)):
    """Pre-sampled light rays and wavelengths as a host program (PROPOSAL, Geant ...) would provide them."""
    g = np.random.default_rng(seed)
    d = g.normal(size=(n, 3)) * 0.6 + np.asarray(axis)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = {"position": (np.asarray(origin) + g.normal(size=(n, 3)) * 0.05).astype(np.float32),
            "direction": d.astype(np.float32),
            "startTime": g.uniform(0.0, 10.0, n).astype(np.float32),
            "contrib": g.uniform(0.5, 2.0, n).astype(np.float32) * 1e3}
    if polarized:
        ref = np.cross(d, g.normal(size=(n, 3)))
        ref /= np.linalg.norm(ref, axis=1, keepdims=True)
        pol = g.normal(size=(n, 3))
        pol /= np.linalg.norm(pol, axis=1, keepdims=True) * g.uniform(1.0, 3.0, (n, 1))
        rays["stokes"] = np.concatenate([np.ones((n, 1)), pol], 1).astype(np.float32)
        rays["polarizationRef"] = ref.astype(np.float32)
    photons = {"wavelength": g.uniform(350.0, 650.0, n).astype(np.float32),
               "contrib": g.uniform(0.8, 1.2, n).astype(np.float32)}
    return rays, photons


def makeLight(kind: str, *, polarized: bool = False, position=(1.0, 0.0, 0.0), direction=(-1.0, 0.0, 0.0)):
    if kind == "spherical":
        return light.SphericalLightSource(position=position, timeRange=(0.0, 10.0), budget=1e5)
    if kind == "cone":
        return light.ConeLightSource(position=position, direction=direction, timeRange=(0.0, 10.0), cosOpeningAngle=0.8,
                                     budget=1e5, polarizationReference=(0.0, 1.0, 0.0), stokes=(1.0, 0.9, 0.1, -0.3),
                                     polarized=polarized)
    if kind == "pencil":
        return light.PencilLightSource(position=position, direction=direction, timeRange=(0.0, 1.0), budget=1e3,
                                       stokes=(1.0, 1.0, 0.0, 0.0), polarizationRef=(0.0, 1.0, 0.0))
    if kind == "cascade":
        return light.ParticleCascadeLightSource(position, 2.0, direction, effectiveLength=5300.0, a_angular=0.54,
                                                b_angular=3.3, a_long=4.6, b_long=0.58)
    if kind == "cherenkov":
        end = tuple(position[k] + 20.0 * direction[k] for k in range(3))
        return light.CherenkovLightSource(trackStart=position, trackEnd=end, startTime=1.0,
                                          endTime=1.0 + 20.0 / 0.299792458, usePhotonCount=True)
    if kind == "cherenkovTrack":
        d = np.asarray(direction, dtype=np.float64)
        side = np.cross(d, (0.3, 0.5, 0.8))
        pts = [np.asarray(position) + 5.0 * k * d + (0.4 * (k % 2)) * side for k in range(6)]
        t = np.concatenate([[0.0], np.cumsum(np.linalg.norm(np.diff(pts, axis=0), axis=1))]) / 0.299792458
        track = light.ParticleTrack(8)
        track.setVertices(np.concatenate([np.array(pts), t[:, None]], 1))
        return light.CherenkovTrackLightSource(track)
    if kind == "muon":
        end = tuple(position[k] + 30.0 * direction[k] for k in range(3))
        return light.MuonTrackLightSource(position, 1.0, end, 1.0 + 30.0 / 0.299792458, muonEnergy=1e3)
    raise ValueError(kind)


def volumeScenario(N: int, *, resp: str = "record", callback: str | None = None, mis: bool = True, direct: bool = True,
                   polarized: bool = False, lightKind: str = "spherical", targetKind: str = "sphere",
                   wavelength: str = "const", water: bool = False, nScattering: int = 10, key: int = 42,
                   scatterCoefficient: float = 0.05, nBins: int = 100, binSize: float = 5.0, maxTime: float = 1000.0,
                   respCapacity: int = 0, photon=None):
    """Config C1 (examples/01 of the reference) and its flag variations (tests/test_trace.py:90-212)."""
    model = WaterTestModel() if water else HomogenousModel()
    med = model.createMedium()
    store = material.MaterialStore([], media=[med])
    if lightKind == "host":
        tgt, src = target.SphereTarget(position=(-1.0, 0.0, 0.0), radius=0.5), None
    elif targetKind == "sphere":
        tgt = target.SphereTarget(position=(-1.0, 0.0, 0.0), radius=0.5)
        src = makeLight(lightKind, polarized=polarized)
    elif targetKind == "flat":
        tgt = target.FlatTarget(width=1.2, length=0.8, position=(-1.0, 0.1, 0.0), direction=(1.0, 0.2, -0.1),
                                up=(0.0, 0.0, 1.0))
        src = makeLight(lightKind, polarized=polarized)
    elif targetKind == "disk":
        tgt = target.DiskTarget(position=(-1.0, 0.0, 0.1), radius=0.6, normal=(0.9, -0.3, 0.2))
        src = makeLight(lightKind, polarized=polarized)
    else:
        tgt = target.InnerSphereTarget(position=(0.0, 0.0, 0.0), radius=8.0)
        src = makeLight(lightKind, polarized=polarized, position=(0.5, 0.2, -0.1), direction=(0.0, 0.6, 0.8))
    wl = light.ConstWavelengthSource(500.0) if wavelength == "const" else light.UniformWavelengthSource(
        lambdaRange=(400.0, 700.0))
    if wavelength == "function":
        wl = light.FunctionWavelengthSource(lambda lam: np.exp(-0.5 * ((lam - 480.0) / 60.0) ** 2) + 0.1,
                                            lambdaRange=(350.0, 650.0), numSamples=256)
    if lightKind == "host":
        # two batches worth of host generated rays / photons, streamed batch by batch
        rays, photons = hostSourceData(2 * N, polarized=polarized)
        src = light.StreamingHostLightSource(N, data=rays, polarized=polarized)
        if wavelength == "host":
            wl = light.StreamingHostWavelengthSource(N, data=photons)
    r = makeResponse(resp, polarized=polarized, nBins=nBins, binSize=binSize, capacity=respCapacity)
    rng = random.PhiloxRNG(key=key)
    cb = makeCallback(callback, N, (photon[0] * photon[1] if photon else nScattering) + 2, polarized)
    if photon:
        # VolumePhotonTracer (examples of the reference's tests/test_trace.py:1605-1690): runs x events per run
        tr = trace.VolumePhotonTracer(N, src, tgt, wl, r, rng, medium=store.media[med.name], nScatteringPerRun=photon[1],
                                      nRuns=photon[0], callback=cb, polarized=polarized, maxTime=maxTime)
    else:
        tr = trace.VolumeForwardTracer(N, src, tgt, wl, r, rng, medium=store.media[med.name], nScattering=nScattering,
                                       scatterCoefficient=scatterCoefficient, callback=cb, polarized=polarized,
                                       disableDirectLighting=not direct, disableTargetSampling=not mis, maxTime=maxTime)
    rng.autoAdvance = tr.nRNGSamples
    return {"tracer": tr, "response": r, "callback": cb, "rng": rng, "store": store}


def _sceneCommon(N, instances, store, medium, src, wl, resp, callback, polarized, maxPathLength, targetId, guide, key,
                 nBins, binSize, **kw):
    sc = scene.Scene(instances, store, medium=medium)
    r = makeResponse(resp, polarized=polarized, nBins=nBins, binSize=binSize, capacity=kw.pop("respCapacity", 0),
                     nObjects=len(instances))
    rng = random.PhiloxRNG(key=key)
    photon = kw.pop("photon", None)
    if photon:  # ScenePhotonTracer: (nRuns, nScatteringPerRun)
        cb = makeCallback(callback, N, photon[0] * photon[1] + 1, polarized)
        tr = trace.ScenePhotonTracer(N, src, wl, r, rng, sc, nRuns=photon[0], nScatteringPerRun=photon[1],
                                     targetId=targetId, callback=cb, polarized=polarized, **kw)
        rng.autoAdvance = tr.nRNGSamples
        return {"tracer": tr, "response": r, "callback": cb, "rng": rng, "store": store, "scene": sc}
    cb = makeCallback(callback, N, maxPathLength + 1, polarized)
    tr = trace.SceneForwardTracer(N, src, wl, r, rng, sc, maxPathLength=maxPathLength, targetId=targetId,
                                  targetGuide=guide, callback=cb, polarized=polarized, **kw)
    rng.autoAdvance = tr.nRNGSamples
    return {"tracer": tr, "response": r, "callback": cb, "rng": rng, "store": store, "scene": sc}


def sceneC2(N: int, *, resp: str = "record", callback: str | None = None, polarized: bool = True,
            lightKind: str = "pencil", key: int = 0xC0FFEE, nBins: int = 100, binSize: float = 2.0, **kw):
    """Config C2: pencil / cone beam into a glass sphere and a detector cube, polarized."""
    water = WaterTestModel().createMedium()
    glass = material.BK7Model().createMedium()
    mats = [material.Material("glassSphere", glass, water, flags="TR"),
            material.Material("detector", None, water, flags="BD")]
    store = material.MaterialStore(mats, media=[water, glass])
    meshes = scene.MeshStore({"sphere": os.path.join(ASSETS, "sphere.stl"), "cube": os.path.join(ASSETS, "cube.ply")})
    inst = [
        meshes.createInstance("sphere", "glassSphere", scene.Transform.TRS(scale=0.8, translate=(3.0, 0.0, 0.0))),
        meshes.createInstance("cube", "detector", scene.Transform.TRS(scale=0.5, translate=(6.0, 0.0, 0.0)),
                              detectorId=1),
    ]
    if lightKind == "pencil":
        src = light.PencilLightSource(position=(0.0, 0.05, 0.02), direction=(1.0, 0.0, 0.0), timeRange=(0.0, 1.0),
                                      stokes=(1.0, 1.0, 0.0, 0.0), polarizationRef=(0.0, 1.0, 0.0), budget=1e3)
    elif lightKind == "host":
        rays, _ = hostSourceData(2 * N, polarized=polarized, origin=(0.0, 0.0, 0.0), axis=(3.0, 0.0, 0.0))
        src = light.StreamingHostLightSource(N, data=rays, polarized=polarized)
    else:
        src = light.ConeLightSource(position=(0.0, 0.0, 0.0), direction=(1.0, 0.0, 0.0), cosOpeningAngle=0.95,
                                    timeRange=(0.0, 1.0), budget=1e3, polarized=polarized,
                                    stokes=(1.0, 0.8, 0.2, 0.1))
    wl = light.UniformWavelengthSource(lambdaRange=(400.0, 700.0))
    return _sceneCommon(N, inst, store, store.media["water"], src, wl, resp, callback, polarized, kw.pop("maxPathLength", 10), 1, None, key,
                        nBins, binSize, **kw)


def sceneC3(N: int, *, resp: str = "record", callback: str | None = None, polarized: bool = False, guide=True,
            key: int = 0xC01DC0FFEE, nBins: int = 200, binSize: float = 1.0, **kw):
    """Config C3: suzanne detector inside a glass torus, scattering medium, target guide."""
    med = HomogenousModel(mu_a=0.05, mu_s=0.05, g=0.9).createMedium()
    glass = material.BK7Model().createMedium()
    mats = [material.Material("det", None, med, flags="BD"), material.Material("glass", glass, med, flags="TR")]
    store = material.MaterialStore(mats, media=[med, glass])
    meshes = scene.MeshStore({"suzanne": os.path.join(ASSETS, "suzanne.stl"),
                              "torus": os.path.join(ASSETS, "torus.stl")})
    inst = [
        meshes.createInstance("suzanne", "det", scene.Transform.TRS(scale=1.0), detectorId=1),
        meshes.createInstance("torus", "glass", scene.Transform.TRS(scale=2.0, rotate=(1.0, 0.0, 0.0, 90.0))),
    ]
    src = light.SphericalLightSource(position=(0.0, -6.0, 0.0), timeRange=(0.0, 5.0), budget=1e5)
    wl = light.ConstWavelengthSource(450.0)
    # guide: True / "sphere", "small", "flat", "disk" (planes facing the light source at y = -6), or False
    if guide == "flat":
        g = target.FlatTargetGuide(width=3.0, height=3.0, position=(0.0, -1.2, 0.0), normal=(0.0, -1.0, 0.0),
                                   up=(0.0, 0.0, 1.0))
    elif guide == "disk":
        g = target.DiskTargetGuide(position=(0.0, -1.2, 0.0), radius=1.6, normal=(0.1, -1.0, 0.05))
    elif guide == "small":
        # much smaller than the detector and off centre: shadow rays outside its cone still reach the detector
        g = target.SphereTargetGuide(position=(0.6, -0.2, 0.3), radius=0.4)
    else:
        g = target.SphereTargetGuide(position=(0.0, 0.0, 0.0), radius=1.5) if guide else None
    return _sceneCommon(N, inst, store, store.media[med.name], src, wl, resp, callback, polarized, 10, 1, g, key,
                        nBins, binSize, **kw)


def sceneC5(N: int, *, resp: str = "hist", callback: str | None = None, polarized: bool = False,
            lattice=(10, 10, 50), pitch=(60.0, 60.0, 10.0), lightKind: str = "cascade", key: int = 0x5EED,
            nBins: int = 500, binSize: float = 10.0, maxPathLength: int = 10, wavelength: str = "uniform", **kw):
    """Config C5: lattice of spherical optical modules (one shared BLAS) in sea water."""
    water = WaterTestModel().createMedium()
    glass = material.BK7Model().createMedium()
    mats = [material.Material("module", glass, water, flags=("DB", "B"))]
    store = material.MaterialStore(mats, media=[water, glass])
    meshes = scene.MeshStore({"sphere": os.path.join(ASSETS, "sphere.stl")})
    nx, ny, nz = lattice
    inst = []
    i = 0
    for ix in range(nx):
        for iy in range(ny):
            for iz in range(nz):
                pos = ((ix - 0.5 * (nx - 1)) * pitch[0], (iy - 0.5 * (ny - 1)) * pitch[1],
                       (iz - 0.5 * (nz - 1)) * pitch[2])
                inst.append(meshes.createInstance("sphere", "module",
                                                  scene.Transform.TRS(scale=0.2159, translate=pos), detectorId=i))
                i += 1
    if lightKind == "cascade":
        src = light.ParticleCascadeLightSource((3.0, 2.0, 1.0), 0.0, (0.0, 0.0, 1.0), effectiveLength=5300.0,
                                               a_angular=0.54, b_angular=3.3, a_long=4.6, b_long=0.58)
    else:
        src = makeLight(lightKind, position=(3.0, 2.0, -100.0), direction=(0.0, 0.0, 1.0))
    wl = light.UniformWavelengthSource(lambdaRange=(300.0, 600.0))
    if wavelength == "host":  # host generated photons ahead of a device light source (no draw for the wavelength)
        wl = light.StreamingHostWavelengthSource(N, data=hostSourceData(2 * N, polarized=False)[1])
    return _sceneCommon(N, inst, store, store.media["water"], src, wl, resp, callback, polarized, maxPathLength, -1,
                        None, key, nBins, binSize, **kw)


# ---------------------------------------------------------------------------------------------
def runGpu(cfg, i: int = 0, *, sync: bool = True):
    """Updates and runs all stages of the tracer once on the GPU (like pl.runPipeline)."""
    import torch
    from theia_b200 import pipeline as pl
    pl.runPipeline(cfg["tracer"].collectStages(), i)
    if sync:
        torch.cuda.synchronize()


def runOracle(cfg, i: int = 0, *, threads: int = 0, oracleScene=None):
    """Runs the same batch on the CPU oracle; results land in the response's host buffers."""
    import oracle
    tr = cfg["tracer"]
    stages = [s for _, s in tr.collectStages()]
    at = next(k for k, s in enumerate(stages) if s is tr)
    for s in stages:
        s.update(i)
    for s in stages[:at]:  # stages ahead of the tracer feed it (host source queues)
        for cmd in s.run(i):
            cmd()
    counters = oracle.trace(tr, i, scene=oracleScene, threads=threads, update=False)
    r = cfg["response"]
    if isinstance(r, response.HistogramHitResponse):  # incl. the kernel histogram
        red = r._reducer
        p = red._block("Params", i)
        parts = red.histIn.numpy().reshape(-1, r.nBins)
        out = oracle.histReduce(parts, p.normalization, clear=True)
        red.histIn.zero_()
        red._buffer[i].numpy()[:] = out
    else:
        for stage in stages[at + 1:]:
            for cmd in stage.run(i):
                cmd()
    return counters


FIELDS = ["position", "direction", "normal", "wavelength", "time", "contrib", "objectId"]
POL_FIELDS = ["stokes", "polarizationRef"]


def sortedTimes(cfg, i: int = 0):
    """Stored (time, objectId) of a StoreTimeHitResponse sorted by (path, event tag)."""
    r = cfg["response"]
    v = r.result(i)
    n = min(v.count, r.capacity)
    path, tag = r.tags()
    path, tag = path[:n].astype(np.int64), tag[:n].astype(np.int64)
    order = np.lexsort((tag, path))
    out = {"path": path[order], "tag": tag[order], "count": v.count, "time": np.array(v["time"][:n])[order]}
    if r.storeObjectId:
        out["objectId"] = np.array(v["objectId"][:n])[order]
    return out


def sortedHits(cfg, i: int = 0):
    """Hits of the last batch sorted by (path, event tag): dict of arrays."""
    r = cfg["response"]
    v = r.queue.view(i)
    n = min(v.count, r.capacity)
    path, tag = r.tags()
    path, tag = path[:n].astype(np.int64), tag[:n].astype(np.int64)
    order = np.lexsort((tag, path))
    out = {"path": path[order], "tag": tag[order], "count": v.count}
    names = FIELDS + (POL_FIELDS if r.polarized else [])
    for f in names:
        out[f] = np.array(v[f][:n])[order]
    return out


def compareHits(a, b, *, rtol: float = 0.0, atol: float = 0.0):
    """Returns (n_common_paths_mismatching, max relative error) between two sortedHits dicts."""
    assert a["count"] == b["count"], f"hit counts differ: {a['count']} vs {b['count']}"
    assert np.array_equal(a["path"], b["path"]) and np.array_equal(a["tag"], b["tag"]), "hit (path, event) tags differ"
    assert np.array_equal(a["objectId"], b["objectId"])
    worst = 0.0
    for f in a:
        if f in ("path", "tag", "count", "objectId"):
            continue
        x, y = a[f].astype(np.float64), b[f].astype(np.float64)
        if rtol == 0.0 and atol == 0.0:
            same = (x == y) | (np.isnan(x) & np.isnan(y))
            assert same.all(), f"field {f}: {np.count_nonzero(~same)} of {same.size} values differ bitwise"
        else:
            err = np.abs(x - y) / (atol + rtol * np.maximum(np.abs(x), np.abs(y)))
            worst = max(worst, float(err.max(initial=0.0)))
            assert (err <= 1.0).all(), f"field {f}: max error ratio {err.max()}"
    return worst


# ---------------------------------------------------------------------------------------------
# analytic scenarios ported from the reference's own tests
class EnergyModel(material.DispersionFreeMedium, material.HenyeyGreensteinPhaseFunction,
                  material.KokhanovskyOceanWaterPhaseMatrix, material.MediumModel):
    """tests/test_trace_energy.py:31-45 of the reference"""

    ModelName = "homogenous"

    def __init__(self, a, s, g, *, n=1.33, ng=1.33) -> None:
        material.DispersionFreeMedium.__init__(self, n=n, ng=ng, mu_a=a, mu_s=s)
        material.HenyeyGreensteinPhaseFunction.__init__(self, g)
        material.KokhanovskyOceanWaterPhaseMatrix.__init__(self, p90=0.66, theta0=0.25, alpha=4.0, xi=25.6)


def beamSplitter(N: int, flag: str, targetId: int, **kw):
    """tests/test_trace.py:1196-1275 of the reference: 45 degree BK7 plate, two detectors."""
    glass = material.BK7Model().createMedium()
    mat = material.Material("mat", glass, None, flags=(flag, "T"))
    det = material.Material("det", glass, None, flags="DB")
    store = material.MaterialStore([mat, det])
    meshes = scene.MeshStore({"cube": os.path.join(ASSETS, "cube.ply")})
    T = scene.Transform
    splitter = meshes.createInstance("cube", "mat", T.TRS(scale=(0.1, 5.0, 5.0), rotate=(0.0, 0.0, 1.0, 45.0),
                                                          translate=(0.141, 0.0, 0.0)))
    refDet = meshes.createInstance("cube", "det", T.TRS(scale=(50.0, 0.5, 50.0), translate=(0.0, -50.0, 0.0)),
                                   detectorId=1)
    transDet = meshes.createInstance("cube", "det", T.TRS(scale=(0.5, 50.0, 50.0), translate=(50.0, 0.0, 0.0)),
                                     detectorId=2)
    sc = scene.Scene([splitter, transDet, refDet], store.material)
    rng = random.PhiloxRNG(key=0xC01DC0FFEE)
    photons = light.UniformWavelengthSource(lambdaRange=(500.0, 500.0))
    src = light.PencilLightSource(position=(-20.0, 0.0, 0.0), direction=(1.0, 0.0, 0.0), timeRange=(0.0, 0.0), budget=1.0)
    rec = response.HitRecorder()
    rec.enableTags()
    tr = trace.SceneForwardTracer(N, src, photons, rec, rng, scene=sc, targetId=targetId, maxPathLength=4, **kw)
    return {"tracer": tr, "response": rec, "callback": None, "rng": rng, "store": store, "scene": sc}


def innerSphereEnergy(N: int, mu_a, mu_s, mu_sample, g, disableDirect, sampleTarget, polarized):
    """tests/test_trace_energy.py:833-938 of the reference: point source inside a spherical target."""
    position, radius = (12.0, 15.0, 0.2), 100.0
    model = EnergyModel(mu_a, mu_s, g)
    med = model.createMedium()
    store = material.MaterialStore([], media=[med])
    rng = random.PhiloxRNG(key=0xC0FFEE)
    photons = light.UniformWavelengthSource(lambdaRange=(400.0, 400.0))
    src = light.SphericalLightSource(position=position, timeRange=(10.0, 10.0), budget=1e9)
    tgt = target.InnerSphereTarget(position=position, radius=radius)
    rec = response.HitRecorder(polarized=polarized)
    rec.enableTags()
    tr = trace.VolumeForwardTracer(N, src, tgt, photons, rec, rng, medium=store.media["homogenous"],
                                   maxTime=float("inf"), nScattering=10, scatterCoefficient=mu_sample,
                                   polarized=polarized, disableDirectLighting=disableDirect,
                                   disableTargetSampling=not sampleTarget)
    rng.autoAdvance = tr.nRNGSamples
    return {"tracer": tr, "response": rec, "callback": None, "rng": rng, "store": store, "model": model}


def runAny(cfg, i: int = 0):
    """Runs one batch on whatever device the scenario was built for."""
    if cfg["tracer"]._device == "cuda":
        runGpu(cfg, i)
    else:
        runOracle(cfg, i)


def hitView(cfg, i: int = 0):
    r = cfg["response"]
    v = r.queue.view(i)
    return v[: min(v.count, r.capacity)]


def volumeC4(N: int, *, key: int = 0xABBA, nBins: int = 100, binSize: float = 5.0):
    """
    Config C4 (examples/03_multiple_lightsources.ipynb cells 4-8 of the reference): several
    VolumeForwardTracer pipelines that share target, wavelength source, response and rng and are
    configured by tasks. Returns the named pipelines, the task list and the shared stages.
    """
    from theia_b200 import pipeline as pl
    med = WaterTestModel().createMedium()
    store = material.MaterialStore([], media=[med])
    rng = random.PhiloxRNG(key=key)
    resp = response.HistogramHitResponse(response.UniformValueResponse(), nBins=nBins, binSize=binSize)
    tgt = target.SphereTarget(position=(-1.0, 0.0, 0.0), radius=0.5)
    wl = light.ConstWavelengthSource(500.0)
    sources = {
        "sphere": light.SphericalLightSource(),
        "pencil": light.PencilLightSource(),
        "cone": light.ConeLightSource(),
        "cascade": makeLight("cascade"),
        "muon": makeLight("muon"),
    }
    tracers = {name: trace.VolumeForwardTracer(N, src, tgt, wl, resp, rng, nScattering=10,
                                               medium=store.media[med.name]) for name, src in sources.items()}
    rng.autoAdvance = max(t.nRNGSamples for t in tracers.values())
    pipelines = {name: pl.Pipeline(t.collectStages()) for name, t in tracers.items()}
    tasks = [
        ("sphere", {"lightSource__budget": 1e5, "lightSource__timeRange": (0.0, 10.0),
                    "lightSource__position": (1.0, 0.0, 0.0)}),
        ("sphere", {}),
        ("pencil", {"lightSource__position": (0.0, 1.0, 0.0), "lightSource__direction": (-0.8, -0.36, -0.24),
                    "lightSource__budget": 5e6, "lightSource__timeRange": (8.0, 8.5)}),
        ("pencil", {}),
        ("cone", {"lightSource__position": (1.0, 0.5, 0.0), "lightSource__direction": (-1.0, 0.0, 0.0),
                  "lightSource__cosOpeningAngle": 0.7, "lightSource__budget": 1e5,
                  "lightSource__timeRange": (0.0, 5.0)}),
        ("cascade", {}),
        ("muon", {}),
        ("sphere", {"lightSource__position": (0.0, 2.0, 0.0)}),
    ]
    return {"pipelines": pipelines, "tracers": tracers, "tasks": tasks, "response": resp, "rng": rng, "store": store}
