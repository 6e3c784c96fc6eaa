"""
Responses beyond record / histogram on the CPU oracle: kernel histogram, store time, object
histogram (sh/response.histogram.kernel.glsl, sh/response.time.store.glsl of the reference and the
per object extension), checked against independent numpy statements over the recorded hits of
the same batch. CPU only; the CUDA side is compared with the oracle in tests/test_gpu_parity.py.
"""

import numpy as np
import pytest
from scipy.stats import norm

import harness as H
import oracle
from theia_b200 import _device, response


def recorded(builder, N, **kw):
    with _device.useDevice("cpu"):
        cfg = builder(N, resp="record", **kw)
        H.runOracle(cfg)
        v = H.hitView(cfg)
        return {f: np.array(v[f]) for f in ("time", "contrib", "objectId")}


def test_kernel_cdf_is_a_normal_cdf():
    # Vazquez-Leal et al. (2011), eq. 4.3 as written in sh/response.histogram.kernel.glsl:62-70
    x = np.linspace(-6.0, 6.0, 2001).astype(np.float32)
    t = oracle.mathEval("tanh", (7.779374467827938 * x - 55.5 * oracle.mathEval("atan", 0.1257926109373887 * x)).astype(np.float32))
    cdf = 0.5 + 0.5 * t.astype(np.float64)
    assert np.abs(cdf - norm.cdf(x.astype(np.float64))).max() < 2e-4


def test_kernel_histogram_on_the_oracle():
    N, nBins, binSize = 1 << 14, 64, 2.0
    hits = recorded(H.volumeScenario, N)
    with _device.useDevice("cpu"):
        cfg = H.volumeScenario(N, resp="kernel", nBins=nBins, binSize=binSize)
        H.runOracle(cfg)
        got = cfg["response"].result(0).astype(np.float64)
        p = cfg["response"].getParams()
    bw, support = p["kernelBandwidth"], p["kernelSupport"]
    assert bw == pytest.approx(1.6) and support == pytest.approx(5.0)

    def cdf(x):
        return 0.5 + 0.5 * np.tanh(7.779374467827938 * x - 55.5 * np.arctan(0.1257926109373887 * x))

    want = np.zeros(nBins)
    t, w = hits["time"].astype(np.float64), hits["contrib"].astype(np.float64)
    first = np.maximum(np.floor((t - support) / binSize), 0).astype(int)
    last = np.minimum(np.ceil((t + support) / binSize), nBins).astype(int) - 1
    for k in range(len(t)):
        edges = np.arange(first[k], last[k] + 2) * binSize - t[k]
        if len(edges) >= 2:
            want[first[k]:last[k] + 1] += np.diff(cdf(edges / bw)) * w[k]
    want /= N
    assert want.sum() > 0
    assert np.allclose(got, want, rtol=2e-3, atol=1e-6 * want.max())


def test_store_time_response_on_the_oracle():
    N = 1 << 14
    with _device.useDevice("cpu"):
        cfg = H.sceneC2(N, resp="time", lightKind="cone")
        cfg["tracer"].source.setParams(budget=0.5)
        cfg["rng"].autoAdvance = 0  # same counter offset in both runs
        r = cfg["response"]
        assert r.nRNGSamples == 1 and r.storeObjectId
        H.runOracle(cfg)
        t = H.sortedTimes(cfg)
        rec = H.sceneC2(N, resp="record", lightKind="cone")
        rec["tracer"].source.setParams(budget=0.5)
        rec["rng"].autoAdvance = 0
        H.runOracle(rec)
        hits = H.sortedHits(rec)
    assert t["count"] > 500 and np.all(t["objectId"] == 1)
    # first hits of a path: same path prefix in both runs (the acceptance draw comes after the hit),
    # so every stored first hit is a recorded first hit with the same time ...
    firstRec = {p: tm for p, tm, tg in zip(hits["path"][::-1], hits["time"][::-1], hits["tag"][::-1])}
    firstT = {}
    for p, tm in zip(t["path"], t["time"]):
        firstT.setdefault(p, tm)
    same = sum(1 for p, tm in firstT.items() if firstRec.get(p) == tm)
    assert same > 0.9 * len(firstT)
    # ... and the accepted fraction is the mean detection probability
    prob = np.minimum(hits["contrib"].astype(np.float64), 1.0)
    assert t["count"] == pytest.approx(prob.sum(), abs=5.0 * np.sqrt(prob.sum()) + 0.03 * prob.sum())
    # times only
    with _device.useDevice("cpu"):
        r2 = response.StoreTimeHitResponse(response.UniformValueResponse(), capacity=16, storeObjectId=False)
        assert r2.queue.view(0).fields == ["time"] and r2.queue.buffer(0).numel() >= 4 + 4 * 16


def test_object_histogram_on_the_oracle():
    N = 1 << 14
    kw = dict(lattice=(4, 4, 4), pitch=(3.0, 3.0, 3.0), nBins=40, binSize=2.5)
    hits = recorded(H.sceneC5, N, **kw)
    with _device.useDevice("cpu"):
        cfg = H.sceneC5(N, resp="objhist", **kw)
        H.runOracle(cfg)
        got = cfg["response"].result(0).astype(np.float64)
        # the device array is cleared for the next batch
        assert float(cfg["response"].tensor.sum()) == 0.0
    bins = np.floor(hits["time"].astype(np.float32) / np.float32(2.5)).astype(np.int64)
    ok = (bins >= 0) & (bins < 40)
    want = np.zeros((64, 40))
    np.add.at(want, (hits["objectId"][ok], bins[ok]), hits["contrib"].astype(np.float64)[ok])
    want /= N
    assert want.sum() > 0 and got.shape == (64, 40)
    assert np.allclose(got, want, rtol=1e-4, atol=1e-7 * want.max())
