"""
GPU parity tests: the CUDA path (through the C-ABI) against the CPU oracle on the same
seeded inputs. Integer / index results must match bit-exact; with the arithmetic contract of
DESIGN.md (same operation order, deterministic transcendental functions) the float results
match bit-exact too, which is far inside the 1e-5 relative tolerance north_star states.
"""

import ctypes

import numpy as np
import pytest
import torch

import harness as H
from theia_b200 import _device, _lib

pytestmark = pytest.mark.gpu

REL_TOL = 1e-5  # tolerance stated by BASELINE.json north_star (float fields of a hit)


def both(builder, N, **kw):
    with _device.useDevice("cuda"):
        g = builder(N, **kw)
    with _device.useDevice("cpu"):
        c = builder(N, **kw)
    return g, c


# ------------------------------------------------------------------ Philox / math
def test_philox_known_answers(cuda, orc):
    ctr = np.array([[0, 0, 0, 0], [0xFFFFFFFF] * 4, [0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344]], np.uint32)
    key = np.array([[0, 0], [0xFFFFFFFF] * 2, [0xA4093822, 0x299F31D0]], np.uint32)
    want = np.array([[0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8], [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD],
                     [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]], np.uint32)
    c = torch.from_numpy(ctr.view(np.int32)).cuda()
    k = torch.from_numpy(key.view(np.int32)).cuda()
    out = torch.zeros(3, 4, dtype=torch.int32, device="cuda")
    _lib.check(_lib.load().theia_philox_raw(c.data_ptr(), k.data_ptr(), 3, out.data_ptr(), 0))
    got = out.cpu().numpy().view(np.uint32)
    assert np.array_equal(got, want)
    assert np.array_equal(orc.philoxRaw(ctr, key), want)


@pytest.mark.parametrize("key,offset", [(0xC0FFEE, 0), (42, (1 << 128) - 5), ((1 << 64) - 1, (1 << 32) - 2)])
def test_philox_stream_bit_exact(cuda, orc, key, offset):
    from theia_b200.random import PhiloxRNG, RNGBufferSink
    streams, samples = 4096, 256
    with _device.useDevice("cuda"):
        rng = PhiloxRNG(key=key, offset=offset)
        sink = RNGBufferSink(rng, streams, samples)
        from theia_b200 import pipeline as pl
        pl.runPipeline([rng, sink])
        torch.cuda.synchronize()
        got = sink.tensor.cpu().numpy().reshape(streams, samples)
    want = orc.philoxFill(key, offset, streams, samples)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert got.min() >= 0.0 and got.max() < 1.0


@pytest.mark.parametrize("sampleDim", [1, 2])
def test_rngSink(cuda, orc, sampleDim):
    # tests/test_random.py:36-64 of the reference: 250 streams x 800 samples from stream 12 / draw 316 on
    from theia_b200 import pipeline as pl
    from theia_b200.random import PhiloxRNG, RNGBufferSink
    with _device.useDevice("cuda"):
        rng = PhiloxRNG(key=0xC0FFEE, offset=77)
        sink = RNGBufferSink(rng, 250, 800, baseStream=12, baseCount=316, sampleDim=sampleDim)
        pl.runPipeline([rng, sink])
        torch.cuda.synchronize()
        got = sink.tensor.cpu().numpy().reshape(sink.shape)
    want = orc.philoxFill(0xC0FFEE, 77, 250, 800, baseStream=12, baseCount=316, dim=sampleDim).reshape(sink.shape)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    if sampleDim == 2:  # random2D_s(stream, i) = (draw i, draw i + 1)
        assert np.array_equal(got[:, 1:, 0], got[:, :-1, 1])


def test_philox_uniformity(cuda):
    # the reference's own pin (tests/test_random.py:67-81): 20 bins, max deviation < 0.5 %
    lib = _lib.load()
    from theia_b200 import _abi
    p = _abi.PhiloxParams()
    p.key[0] = 0xC0FFEE
    streams, samples = 4096, 8192
    out = torch.zeros(streams * samples, dtype=torch.float32, device="cuda")
    _lib.check(lib.theia_philox_fill(ctypes.byref(p), 0, 0, streams, samples, 1, out.data_ptr(), 0))
    hist = torch.histc(out, bins=20, min=0.0, max=1.0).cpu().numpy() / out.numel()
    assert np.abs(hist * 20 - 1.0).max() < 5e-3


@pytest.mark.parametrize("fn", ["sin", "cos", "exp", "log", "pow", "tanh", "atan"])
def test_device_math_bit_exact(cuda, orc, fn):
    rng = np.random.default_rng(7)
    n = 1 << 18
    if fn in ("sin", "cos", "tanh"):
        x = rng.uniform(-20.0, 20.0, n)
    elif fn == "atan":
        x = np.tan(rng.uniform(-1.57, 1.57, n))
    elif fn == "exp":
        x = rng.uniform(-110.0, 90.0, n)
    elif fn == "log":
        x = np.exp(rng.uniform(-100.0, 80.0, n))
    else:
        x = np.exp(rng.uniform(-20.0, 5.0, n))
    x = x.astype(np.float32)
    y = rng.uniform(-4.0, 4.0, n).astype(np.float32)
    xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    out = torch.zeros(n, dtype=torch.float32, device="cuda")
    import oracle
    _lib.check(_lib.load().theia_math_eval(oracle.MATH_FN[fn], xd.data_ptr(), yd.data_ptr(), n, out.data_ptr(), 0))
    got = out.cpu().numpy()
    want = orc.mathEval(fn, x, y)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    ref = {"sin": np.sin, "cos": np.cos, "exp": np.exp, "log": np.log, "tanh": np.tanh, "atan": np.arctan}.get(fn)
    exact = ref(x.astype(np.float64)) if ref else np.power(x.astype(np.float64), y.astype(np.float64))
    ok = np.isfinite(exact) & (np.abs(exact) > 1e-30) & (np.abs(exact) < 1e30)
    # sin, cos, tanh: absolute accuracy near their zeros
    rel = np.abs(got[ok] - exact[ok]) / np.maximum(np.abs(exact[ok]), 1.0 if fn == "tanh" else (1e-6 if fn in ("sin", "cos") else 0.0))
    assert rel.max() < 4e-7  # a few ulp, inside what GLSL guarantees for its built-ins


# ------------------------------------------------------------------ ray query
def _randomRays(rng, n, lo, hi, tmax):
    o = rng.uniform(lo, hi, (n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((n, 8), np.float32)
    rays[:, :3] = o
    rays[:, 3] = tmax
    rays[:, 4:7] = d
    return rays


@pytest.mark.parametrize("queryKernel", [0, 1, 2], ids=["direct", "phased", "queued"])
@pytest.mark.parametrize("builder,kw,box", [
    (H.sceneC2, {}, (-2.0, 8.0)),
    (H.sceneC3, {}, (-4.0, 4.0)),
    (H.sceneC5, {"lattice": (6, 6, 6), "pitch": (2.0, 2.0, 2.0), "resp": "record"}, (-8.0, 8.0)),
])
def test_ray_query_matches_brute_force(cuda, orc, builder, kw, box, queryKernel):
    g, c = both(builder, 256, **kw)
    rng = np.random.default_rng(3)
    n = 200000
    rays = _randomRays(rng, n, box[0], box[1], 1000.0)
    # axis aligned and grazing rays
    rays[:1000, 4:7] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, 1000)] * rng.choice([-1.0, 1.0], (1000, 1))
    brute = orc.OracleScene(c["scene"], bruteForce=True)
    ids_w, tuv_w = brute.query(rays)
    rd = torch.from_numpy(rays).cuda()
    ids = torch.zeros(n, 4, dtype=torch.int32, device="cuda")
    tuv = torch.zeros(n, 4, dtype=torch.float32, device="cuda")
    st = torch.zeros(4, dtype=torch.int64, device="cuda")
    lib = _lib.load()
    # the same queries through the one-thread-per-ray traversal and through both wavefront ray query kernels
    _lib.check(lib.theia_set_option(b"query_kernel", queryKernel))
    try:
        _lib.check(lib.theia_tlas_query(g["scene"].tlas.data_ptr(), rd.data_ptr(), n, ids.data_ptr(), tuv.data_ptr(),
                                        st.data_ptr(), 0))
        torch.cuda.synchronize()
    finally:
        lib.theia_set_option(b"query_kernel", 0)
    ids_g, tuv_g = ids.cpu().numpy(), tuv.cpu().numpy()
    assert np.count_nonzero(ids_w[:, 0] >= 0) > n // 100
    assert np.array_equal(ids_g, ids_w), f"{np.count_nonzero((ids_g != ids_w).any(1))} rays hit a different triangle"
    assert np.array_equal(tuv_g.view(np.uint32), tuv_w.view(np.uint32))
    stats = st.cpu().numpy()
    assert queryKernel != 0 or (stats[3] == n and stats[0] > 0)


def test_sweep_builder_option_gives_the_same_ray_queries(cuda):
    # THEIA_BVH_BUILDER=sah (read once per process): the brute force comparison above in a child process
    import os, subprocess, sys
    if os.environ.get("THEIA_BVH_BUILDER") == "sah":
        pytest.skip("already inside the child")
    env = dict(os.environ, THEIA_BVH_BUILDER="sah")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-q", "-m", "gpu", "-x", "-k",
                        "brute_force or (scene_tracer_hits and sceneC3 and wavefront)"], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert " passed" in r.stdout


# ------------------------------------------------------------------ tracers, per path parity
VOLUME_CASES = [
    dict(),
    dict(mis=False),
    dict(mis=False, direct=False),
    dict(direct=False),
    dict(polarized=True, water=True),
    dict(polarized=True, water=True, lightKind="cone", mis=False),
    dict(lightKind="pencil", polarized=True, water=True),
    dict(lightKind="cascade", water=True, wavelength="uniform"),
    dict(lightKind="muon", water=True, wavelength="uniform"),
    dict(targetKind="inner", mis=True),
    dict(targetKind="inner", mis=False, polarized=True, water=True),
    dict(lightKind="cherenkov", water=True, wavelength="uniform"),
    dict(lightKind="cherenkov", water=True, wavelength="function", polarized=True),
    dict(lightKind="cherenkovTrack", water=True, wavelength="function", mis=False),
    dict(lightKind="cherenkovTrack", water=True, wavelength="uniform", polarized=True),
    dict(lightKind="host"),
    dict(lightKind="host", wavelength="host", polarized=True, water=True, mis=False),
    dict(targetKind="flat"),
    dict(targetKind="flat", mis=False, polarized=True, water=True, lightKind="cone"),
    dict(targetKind="disk"),
    dict(targetKind="disk", direct=False, polarized=True, water=True),
]


@pytest.mark.parametrize("kw", VOLUME_CASES, ids=lambda k: ",".join(f"{a}={b}" for a, b in k.items()) or "C1")
def test_volume_tracer_hits_bit_exact(cuda, orc, kw):
    N = 1 << 14
    g, c = both(H.volumeScenario, N, resp="record", **kw)
    H.runGpu(g)
    H.runOracle(c)
    a, b = H.sortedHits(g), H.sortedHits(c)
    assert b["count"] > 30
    H.compareHits(a, b)


SCENE_CASES = [
    (H.sceneC2, dict()),
    (H.sceneC2, dict(lightKind="cone")),
    (H.sceneC2, dict(polarized=False, lightKind="cone", useRefractedHitDir=True)),
    (H.sceneC2, dict(lightKind="cone", disableTransmission=True)),
    (H.sceneC3, dict()),
    (H.sceneC3, dict(guide=False)),
    (H.sceneC3, dict(polarized=True, disableDirectLighting=True)),
    (H.sceneC2, dict(lightKind="host")),
    (H.sceneC3, dict(guide="flat")),
    (H.sceneC3, dict(guide="disk", polarized=True)),
    (H.sceneC3, dict(guide="small")),
    (H.sceneC5, dict(resp="record", lattice=(6, 6, 6), pitch=(4.0, 4.0, 4.0))),
    # host wavelength source (no draw) ahead of the cascade light, whose gamma variate the wavefront pregenerates
    (H.sceneC5, dict(resp="record", lattice=(6, 6, 6), pitch=(4.0, 4.0, 4.0), wavelength="host")),
]


@pytest.mark.parametrize("mode", ["single", "wavefront", "wavefront-simple", "wavefront-queued"])
@pytest.mark.parametrize("builder,kw", SCENE_CASES, ids=lambda v: getattr(v, "__name__", None) or ",".join(
    f"{a}={b}" for a, b in v.items()) or "default")
def test_scene_tracer_hits_bit_exact(cuda, orc, builder, kw, mode):
    N = 1 << 14
    kw = dict(kw)
    kw.setdefault("resp", "record")
    kw["mode"] = mode.split("-")[0]
    # both ray query kernels of the wavefront tracer (normally picked by timing) are covered
    lib = _lib.load()
    forced = {"simple": 1, "queued": 2}.get(mode.split("-")[-1], 0)
    lib.theia_set_option(b"wf_trace_kernel", forced)
    lib.theia_set_option(b"wf_shadow_trace_kernel", forced)
    try:
        g, c = both(builder, N, **kw)
        H.runGpu(g)
    finally:
        lib.theia_set_option(b"wf_trace_kernel", -1)
        lib.theia_set_option(b"wf_shadow_trace_kernel", -1)
    H.runOracle(c)
    a, b = H.sortedHits(g), H.sortedHits(c)
    assert b["count"] > 0
    H.compareHits(a, b)


@pytest.mark.parametrize("kw", [dict(), dict(guide="small"), dict(guide="flat", polarized=True)],
                         ids=lambda k: ",".join(f"{a}={b}" for a, b in k.items()) or "C3")
def test_shadow_ray_culling_changes_nothing(cuda, kw):
    """Shadow rays that cannot reach the target are not traced (wfTargetBoxKernel): same hits, same counts as with
    every one of them traced."""
    lib = _lib.load()
    runs = []
    for cull in (1, 0):
        lib.theia_set_option(b"wf_shadow_cull", cull)
        try:
            with _device.useDevice("cuda"):
                g = H.sceneC3(1 << 16, resp="record", **kw)
            counters = g["tracer"].enableWorkCounters()
            H.runGpu(g)
            runs.append((H.sortedHits(g), counters.cpu().numpy().copy()))
        finally:
            lib.theia_set_option(b"wf_shadow_cull", 1)
    (a, ca), (b, cb) = runs
    assert a["count"] > 100
    H.compareHits(a, b)
    assert np.array_equal(ca, cb)


@pytest.mark.parametrize("builder,kw", [(H.volumeScenario, dict()), (H.volumeScenario, dict(mis=False)),
                                        (H.sceneC2, dict(lightKind="cone")), (H.sceneC3, dict()),
                                        (H.sceneC2, dict(lightKind="cone", mode="wavefront")),
                                        (H.sceneC3, dict(mode="wavefront"))])
def test_event_statistics_match(cuda, orc, builder, kw):
    N = 1 << 14
    g, c = both(builder, N, resp="empty", callback="stat", **kw)
    H.runGpu(g)
    H.runOracle(c)
    got, want = g["callback"]._read(), c["callback"]._read()
    assert np.array_equal(got, want), f"{got} != {want}"
    assert got[0] == N


@pytest.mark.parametrize("builder,kw", [(H.volumeScenario, dict(polarized=True, water=True)), (H.sceneC2, dict()),
                                        (H.sceneC2, dict(mode="wavefront")), (H.sceneC3, dict(mode="wavefront"))])
def test_track_record_matches(cuda, orc, builder, kw):
    N = 2048
    g, c = both(builder, N, resp="empty", callback="track", **kw)
    H.runGpu(g)
    H.runOracle(c)
    tg, lg, cg = g["callback"].result(0)
    tc, lc, cc = c["callback"].result(0)
    assert np.array_equal(lg, lc) and np.array_equal(cg, cc)
    for k in range(N):
        n = min(int(lg[k]) + 1, tg.shape[1])
        assert np.array_equal(tg[k, :n].view(np.uint32), tc[k, :n].view(np.uint32)), f"track {k}"


@pytest.mark.parametrize("builder,kw", [(H.volumeScenario, dict()), (H.sceneC3, dict(mode="single")),
                                        (H.sceneC3, dict(mode="wavefront")),
                                        (H.sceneC2, dict(lightKind="cone", polarized=False))])
def test_histogram_matches_oracle(cuda, orc, builder, kw):
    N = 1 << 16
    g, c = both(builder, N, resp="hist", **kw)
    H.runGpu(g)
    H.runOracle(c)
    got, want = g["response"].result(0).astype(np.float64), c["response"].result(0).astype(np.float64)
    assert want.sum() > 0
    # same hits, different float summation order (the reference compares with rtol 1e-4:
    # tests/test_response.py:92-99)
    assert np.allclose(got, want, rtol=1e-4, atol=1e-7 * want.max())


# ------------------------------------------------------------------ sharded run / profiler
@pytest.mark.parametrize("builder,kw", [(H.volumeScenario, dict()), (H.sceneC2, dict(lightKind="cone")),
                                        (H.sceneC3, dict(mode="wavefront"))],
                         ids=["volume", "c2", "c3"])
def test_kernel_histogram_matches_oracle(cuda, orc, builder, kw):
    # sh/response.histogram.kernel.glsl: same bins, same weights; only the order of the float sums differs
    N = 1 << 15
    g, c = both(builder, N, resp="kernel", nBins=64, binSize=2.0, **kw)
    H.runGpu(g)
    H.runOracle(c)
    a, b = g["response"].result(0).astype(np.float64), c["response"].result(0).astype(np.float64)
    assert b.sum() > 0
    assert np.allclose(a, b, rtol=2e-4, atol=1e-6 * b.max())
    # smearing conserves what lands inside the histogram: compare with the plain histogram away from the edges
    g2, _ = both(builder, N, resp="hist", nBins=64, binSize=2.0, **kw)
    H.runGpu(g2)
    plain = g2["response"].result(0).astype(np.float64)
    inner = slice(4, 60)
    assert a[inner].sum() == pytest.approx(plain[inner].sum(), rel=0.15)


@pytest.mark.parametrize("builder,kw", [(H.volumeScenario, dict()), (H.volumeScenario, dict(mis=False, polarized=True, water=True)),
                                        (H.sceneC2, dict(lightKind="cone", mode="single")),
                                        (H.sceneC2, dict(lightKind="cone", mode="wavefront")),
                                        (H.sceneC3, dict(mode="single")), (H.sceneC3, dict(mode="wavefront"))],
                         ids=["volume", "volume-pol", "c2-single", "c2-wavefront", "c3-single", "c3-wavefront"])
def test_store_time_response_bit_exact(cuda, orc, builder, kw):
    # sh/response.time.store.glsl: the acceptance draw comes from the path's own stream, so the set of stored
    # photons, their times and object ids AND all later draws of the path (the following hits) must match
    N = 1 << 14
    g, c = both(builder, N, resp="time", **kw)
    # detection probabilities in (0, 1): a budget below one, so that the draw decides
    for cfg in (g, c):
        cfg["tracer"].source.setParams(budget=0.5)
    H.runGpu(g)
    H.runOracle(c)
    a, b = H.sortedTimes(g), H.sortedTimes(c)
    assert b["count"] > 20 and a["count"] == b["count"]
    for f in ("path", "tag", "time", "objectId"):
        assert np.array_equal(a[f], b[f]), f


def test_object_histogram_matches_recorded_hits(cuda, orc):
    # B200 extension: [object][time] light curves == binning the recorded hits by (objectId, time) on the host
    N = 1 << 16
    kw = dict(lattice=(4, 4, 4), pitch=(3.0, 3.0, 3.0), nBins=40, binSize=2.5)
    with _device.useDevice("cuda"):
        g = H.sceneC5(N, resp="objhist", **kw)
        rec = H.sceneC5(N, resp="record", **kw)
    with _device.useDevice("cpu"):
        c = H.sceneC5(N, resp="objhist", **kw)
    H.runGpu(g)
    H.runGpu(rec)
    H.runOracle(c)
    hist = g["response"].result(0).astype(np.float64)
    assert hist.shape == (64, 40) and hist.sum() > 0
    v = H.hitView(rec)
    bins = np.floor(v["time"].astype(np.float32) / np.float32(2.5)).astype(np.int64)
    ok = (bins >= 0) & (bins < 40)
    want = np.zeros((64, 40))
    np.add.at(want, (np.array(v["objectId"])[ok], bins[ok]), np.array(v["contrib"], dtype=np.float64)[ok])
    want /= N  # the tracer's normalisation
    assert np.allclose(hist, want, rtol=1e-4, atol=1e-7 * want.max())
    assert np.allclose(c["response"].result(0), hist, rtol=1e-4, atol=1e-7 * want.max())
    assert np.count_nonzero(hist.sum(1)) > 16  # many modules saw light


PHOTON_CASES = [
    (H.volumeScenario, dict(photon=(3, 4))),
    (H.volumeScenario, dict(photon=(2, 5), polarized=True, water=True, lightKind="cone")),
    (H.volumeScenario, dict(photon=(1, 12), targetKind="inner", wavelength="uniform", water=True)),
    (H.sceneC2, dict(photon=(3, 4), lightKind="cone", mode="single")),
    (H.sceneC2, dict(photon=(3, 4), lightKind="cone", mode="wavefront")),
    (H.sceneC2, dict(photon=(2, 3), lightKind="cone", polarized=False, mode="wavefront", useRefractedHitDir=True)),
    (H.sceneC3, dict(photon=(4, 3), guide=False, mode="single")),
    (H.sceneC3, dict(photon=(4, 3), guide=False, mode="wavefront")),
]


@pytest.mark.parametrize("resp", ["record", "time"])
@pytest.mark.parametrize("builder,kw", PHOTON_CASES, ids=lambda v: getattr(v, "__name__", None) or ",".join(
    f"{a}={b}" for a, b in v.items()))
def test_photon_tracers_bit_exact(cuda, orc, builder, kw, resp):
    # VolumePhotonTracer / ScenePhotonTracer incl. the draw counter restart at every run boundary;
    # with the store-time response the stored photons are the detected ones
    N = 1 << 15
    g, c = both(builder, N, resp=resp, **kw)
    H.runGpu(g)
    H.runOracle(c)
    if resp == "record":
        a, b = H.sortedHits(g), H.sortedHits(c)
        assert b["count"] > 20
        H.compareHits(a, b)
    else:
        a, b = H.sortedTimes(g), H.sortedTimes(c)
        assert b["count"] > 20 and a["count"] == b["count"]
        for f in ("path", "tag", "time", "objectId"):
            assert np.array_equal(a[f], b[f]), f


@pytest.mark.parametrize("builder,kw", [(H.volumeScenario, dict(photon=(2, 3), polarized=True, water=True)),
                                        (H.sceneC2, dict(photon=(2, 3), lightKind="cone", mode="wavefront")),
                                        (H.sceneC2, dict(photon=(2, 3), lightKind="cone", mode="single"))],
                         ids=["volume", "c2-wavefront", "c2-single"])
def test_photon_tracers_events_match(cuda, orc, builder, kw):
    N = 1 << 12
    for cb in ("stat", "track"):
        g, c = both(builder, N, resp="empty", callback=cb, **kw)
        H.runGpu(g)
        H.runOracle(c)
        if cb == "stat":
            got, want = g["callback"]._read(), c["callback"]._read()
            assert np.array_equal(got, want), f"{got} != {want}"
            assert got[0] == N
        else:
            (ta, la, ca), (tb, lb, cbb) = g["callback"].result(0), c["callback"].result(0)
            assert np.array_equal(la, lb) and np.array_equal(ca, cbb)
            for k in range(N):
                n = min(int(la[k]) + 1, ta.shape[1])
                assert np.array_equal(ta[k, :n].view(np.uint32), tb[k, :n].view(np.uint32)), f"track {k}"


def test_sharded_run_single_rank_matches_oracle(cuda, orc):
    from theia_b200 import distributed
    N, nBatches = 8192, 3
    g, c = both(H.volumeScenario, N, resp="hist", callback="stat")
    res = distributed.ShardedRun(g["tracer"], offset0=11).run(nBatches)
    want = distributed.ShardedRun(c["tracer"], offset0=11, launcher=lambda i: H.runOracle(c, i)).run(nBatches)
    assert res.localBatches == [0, 1, 2]
    assert res.counters == want.counters
    assert res.counters["created"] == N * nBatches
    assert np.allclose(res.lightCurve, want.lightCurve, rtol=1e-5, atol=1e-12)
    # per batch hook: handed over one batch behind the launches, in order, each batch's own curve
    seenG, seenC = [], []
    nB = 5
    resG = distributed.ShardedRun(g["tracer"], offset0=11).run(nB, processFn=lambda b, h: seenG.append((b, np.array(h))))
    distributed.ShardedRun(c["tracer"], offset0=11, launcher=lambda i: H.runOracle(c, i)).run(
        nB, processFn=lambda b, h: seenC.append((b, np.array(h))))
    assert [b for b, _ in seenG] == list(range(nB))
    for (_, hg), (_, hc) in zip(seenG, seenC):
        assert hc.sum() > 0 and np.allclose(hg, hc, rtol=1e-5, atol=1e-12)
    assert np.allclose(np.mean([h for _, h in seenG], axis=0), resG.lightCurve, rtol=1e-5)


def test_sharded_run_c4_tasks_and_gathered_hits_match_oracle(cuda, orc):
    # the product's multi GPU API on the device (one rank): tasks of several pipelines (C4), per object light curves
    # (C5) accumulated without a host round trip, and hits gathered in batch order
    from theia_b200 import distributed
    N = 8192
    g, c = both(H.volumeC4, N)
    res = distributed.ShardedRun(g["tracers"], offset0=5).runTasks(g["tasks"])

    def launcherOf(name):
        return lambda i: H.runOracle({"tracer": c["tracers"][name], "response": c["response"]}, i)

    want = distributed.ShardedRun(c["tracers"], offset0=5, launcher={n: launcherOf(n) for n in c["tracers"]}).runTasks(c["tasks"])
    assert res.taskCounts == want.taskCounts
    for name, curve in want.lightCurves.items():
        assert curve.sum() > 0
        assert np.allclose(res.lightCurves[name], curve, rtol=2e-4, atol=1e-9 * curve.max()), name

    g, c = both(H.sceneC5, N, resp="objhist", lattice=(3, 3, 3), pitch=(3.0, 3.0, 3.0), nBins=40, binSize=2.5)
    res = distributed.ShardedRun(g["tracer"], offset0=2).run(3)
    want = distributed.ShardedRun(c["tracer"], offset0=2, launcher=lambda i: H.runOracle(c, i)).run(3)
    assert want.lightCurve.shape == (27, 40) and want.lightCurve.sum() > 0
    assert np.allclose(res.lightCurve, want.lightCurve, rtol=1e-4, atol=1e-9 * want.lightCurve.max())

    g, c = both(H.volumeScenario, N, resp="record", polarized=True)
    runG = distributed.ShardedRun(g["tracer"], offset0=9)
    runC = distributed.ShardedRun(c["tracer"], offset0=9, launcher=lambda i: H.runOracle(c, i))
    hg, hc = runG.gatherHits(runG.run(3, keepHits=True)), runC.gatherHits(runC.run(3, keepHits=True))
    assert np.array_equal(hg["batch"], hc["batch"]) and len(hc["batch"]) > 0
    for b in range(3):  # within a batch the device records hits in any order: compare as sorted sets
        sel = hc["batch"] == b
        for f in ("time", "contrib", "wavelength"):
            assert np.array_equal(np.sort(hg[f][sel]), np.sort(hc[f][sel])), (b, f)


def test_profile_counters(cuda):
    lib = _lib.load()
    with _device.useDevice("cuda"):
        cfg = H.sceneC2(1 << 16, resp="hist", lightKind="cone", mode="wavefront")
    _lib.check(lib.theia_profile_enable(1))
    H.runGpu(cfg)
    torch.cuda.synchronize()
    ms = (ctypes.c_double * 8)()
    n = (ctypes.c_uint64 * 8)()
    assert lib.theia_profile_read(ms, n, 8) == 8
    _lib.check(lib.theia_profile_enable(0))
    names = [lib.theia_profile_class_name(k).decode() for k in range(8)]
    got = {names[k]: (int(n[k]), float(ms[k])) for k in range(8) if n[k]}
    assert got["wfGenerateKernel"][0] == 1
    assert got["wfTraceKernel"][0] == 10 and got["wfShadeKernel"][0] == 10
    assert all(v[1] > 0.0 for v in got.values())
    H.runGpu(cfg)  # disabled again: nothing accumulates
    assert lib.theia_profile_read(ms, n, 8) == 8 and int(n[3]) == 10


# ------------------------------------------------------------------ C4: several pipelines, shared stages, tasks
def test_c4_multiple_light_sources_scheduler_matches_oracle(cuda, orc):
    # examples/03_multiple_lightsources.ipynb: tasks dealt to named pipelines that share the
    # target, response and rng; every batch must equal the oracle run of the same task
    from theia_b200 import pipeline as pl
    N = 16384
    with _device.useDevice("cuda"):
        g = H.volumeC4(N)
    with _device.useDevice("cpu"):
        c = H.volumeC4(N)
    hists = []

    def process(config, batch, args):
        hists.append(g["response"].result(config).copy())

    sched = pl.PipelineScheduler(g["pipelines"], processFn=process)
    sched.schedule(g["tasks"])
    sched.wait()
    sched.destroy()
    assert len(hists) == len(g["tasks"])
    for n, (name, params) in enumerate(c["tasks"]):
        i = n % 2
        c["pipelines"][name].setParams(**params)
        cfg = {"tracer": c["tracers"][name], "response": c["response"]}
        H.runOracle(cfg, i)
        want = c["response"].result(i)
        assert want.sum() > 0, name
        assert np.allclose(hists[n], want, rtol=2e-4, atol=1e-9 * want.max()), (n, name)
    assert g["rng"].getParam("offset") == c["rng"].getParam("offset") == len(g["tasks"]) * g["rng"].autoAdvance


def test_trace_kernel_is_picked_by_timing(cuda):
    # the first nine batches time the three ray query kernels; afterwards a choice is reported and
    # the results of all batches are those of a run with a forced kernel
    lib = _lib.load()
    N = 1 << 17
    with _device.useDevice("cuda"):
        a = H.sceneC2(N, resp="hist", lightKind="cone", mode="wavefront")
        b = H.sceneC2(N, resp="hist", lightKind="cone", mode="wavefront")
    got, want = [], []
    for k in range(12):
        H.runGpu(a, k % 2)
        torch.cuda.synchronize()
        got.append(a["response"].result(k % 2).copy())
    lib.theia_set_option(b"wf_trace_kernel", 0)
    try:
        for k in range(12):
            H.runGpu(b, k % 2)
            torch.cuda.synchronize()
            want.append(b["response"].result(k % 2).copy())
    finally:
        lib.theia_set_option(b"wf_trace_kernel", -1)
    p, s = ctypes.c_int(-2), ctypes.c_int(-2)
    _lib.check(lib.theia_get_trace_choice(a["scene"].tlas.data_ptr(), N, 0, ctypes.byref(p), ctypes.byref(s)))
    assert p.value in (0, 1, 2) and s.value == 0
    for x, y in zip(got, want):
        assert x.sum() > 0 and np.allclose(x, y, rtol=1e-4, atol=1e-9 * y.max())
