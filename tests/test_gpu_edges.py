"""
Edge cases and statistical agreement of the CUDA path (through the C-ABI) against the oracle:
ragged and empty batches, a hit queue that overflows, the shortest paths, and the chi-square
test on light curves from independent random streams that BASELINE.json's north_star asks for.
"""

import numpy as np
import pytest
import torch
from scipy import stats

import harness as H
from theia_b200 import _device

pytestmark = pytest.mark.gpu


def both(builder, N, **kw):
    with _device.useDevice("cuda"):
        g = builder(N, **kw)
    with _device.useDevice("cpu"):
        c = builder(N, **kw)
    return g, c


@pytest.mark.parametrize("builder,kw", [(H.volumeScenario, dict()), (H.sceneC2, dict(lightKind="cone", mode="single")),
                                        (H.sceneC2, dict(lightKind="cone", mode="wavefront")),
                                        (H.sceneC3, dict(mode="wavefront")),
                                        (H.sceneC3, dict(mode="wavefront", guide="disk", polarized=True)),
                                        (H.sceneC2, dict(lightKind="cone", mode="wavefront", photon=(2, 3)))],
                         ids=["volume", "c2-single", "c2-wavefront", "c3-guide", "c3-disk-guide", "c2-photon"])
@pytest.mark.parametrize("batch", [1, 31, 33, 127, 1000, 4097])
def test_ragged_batches(cuda, orc, builder, kw, batch):
    # batchSize below capacity and not a multiple of the warp / CTA size (src/theia/trace.py:399-408)
    g, c = both(builder, 8192, resp="record", **kw)
    for cfg in (g, c):
        cfg["tracer"].batchSize = batch
    H.runGpu(g)
    H.runOracle(c)
    a, b = H.sortedHits(g), H.sortedHits(c)
    assert a["count"] == b["count"]
    assert (b["path"] < batch).all()
    H.compareHits(a, b)


def test_empty_batch(cuda):
    # the reference accepts batchSize = 0 in the setter and then fails on 1 / batchSize when the
    # stage is updated (src/theia/trace.py:399-408, 431-436): mirrored. At the C-ABI an empty
    # launch is a no-op.
    import ctypes
    from theia_b200 import _lib
    with _device.useDevice("cuda"):
        g = H.volumeScenario(1024, resp="record")
    tr = g["tracer"]
    H.runGpu(g)
    torch.cuda.synchronize()
    before = g["response"].queue.view(0).count
    assert before > 0
    lib = _lib.load()
    L, keep = tr._launch(0)
    L.batch_size = 0
    n0 = lib.theia_kernel_launches()
    _lib.check(lib.theia_trace_forward(ctypes.byref(L)))
    torch.cuda.synchronize()
    assert lib.theia_kernel_launches() == n0
    tr.batchSize = 0
    with pytest.raises(ZeroDivisionError):
        tr.update(0)


@pytest.mark.parametrize("builder,kw", [(H.volumeScenario, dict(targetKind="inner")),
                                        (H.sceneC2, dict(lightKind="pencil", mode="wavefront"))])
def test_hit_queue_overflow_keeps_counting(cuda, orc, builder, kw):
    # sh/response.record.glsl:6-27 + sh/response.queue.glsl: the counter keeps counting, stores beyond
    # the capacity are dropped; what is stored are hits of the run with room for all of them
    N, cap = 4096, 256
    g, c = both(builder, N, resp="record", **kw)
    with _device.useDevice("cuda"):
        small = builder(N, resp="record", respCapacity=cap, **kw)
    assert small["response"].capacity == cap
    H.runGpu(g)
    H.runGpu(small)
    H.runOracle(c)
    full = H.sortedHits(c)
    assert full["count"] > cap
    v = small["response"].queue.view(0)
    assert v.count == full["count"] == g["response"].queue.view(0).count
    stored = np.array(v["time"][:cap])
    assert np.isin(stored.view(np.uint32), full["time"].view(np.uint32)).all()


@pytest.mark.parametrize("builder,kw", [(H.sceneC2, dict(lightKind="cone", maxPathLength=1)),
                                        (H.sceneC2, dict(lightKind="cone", maxPathLength=2)),
                                        (H.volumeScenario, dict(nScattering=1)),
                                        (H.volumeScenario, dict(nScattering=1, mis=False))])
def test_shortest_paths(cuda, orc, builder, kw):
    g, c = both(builder, 4096, resp="record", callback=None, **kw)
    H.runGpu(g)
    H.runOracle(c)
    a, b = H.sortedHits(g), H.sortedHits(c)
    assert a["count"] == b["count"]
    H.compareHits(a, b)


def batchMeans(cfg, nBatches, run):
    out = []
    for k in range(nBatches):
        run(cfg, k % 2)
        out.append(np.array(cfg["response"].result(k % 2), dtype=np.float64))
    return np.stack(out)


@pytest.mark.parametrize("builder,kw", [(H.volumeScenario, dict()), (H.sceneC2, dict(lightKind="cone")),
                                        (H.sceneC3, dict())])
def test_light_curves_agree_statistically(cuda, orc, builder, kw):
    # independent Philox keys on the two sides: chi-square over the bins with per-bin variances from
    # batch means must give p > 0.01 (north_star); and must FAIL for a deliberately shifted curve
    N, nBatches = 1 << 16, 16
    with _device.useDevice("cuda"):
        g = builder(N, resp="hist", key=0x1234, **kw)
    with _device.useDevice("cpu"):
        c = builder(N, resp="hist", key=0x9876, **kw)

    def runGpu(cfg, i):
        H.runGpu(cfg, i)
        torch.cuda.synchronize()

    a = batchMeans(g, nBatches, runGpu)
    b = batchMeans(c, nBatches, H.runOracle)
    ma, mb = a.mean(0), b.mean(0)
    va, vb = a.var(0, ddof=1) / nBatches, b.var(0, ddof=1) / nBatches
    ok = (a > 0).all(0) & (b > 0).all(0)  # bins that every batch populates on both sides
    assert ok.sum() >= 5
    chi2 = float(((ma - mb)[ok] ** 2 / (va + vb)[ok]).sum())
    # batch-mean variances are themselves estimates: use the F-like correction (Hotelling) conservatively
    p = float(stats.chi2.sf(chi2 * (nBatches - 3) / (nBatches - 1), int(ok.sum())))
    assert p > 0.01, (chi2, int(ok.sum()), p)
    shifted = np.roll(mb, 1)
    chi2s = float(((ma - shifted)[ok] ** 2 / (va + vb)[ok]).sum())
    assert stats.chi2.sf(chi2s, int(ok.sum())) < 1e-6
