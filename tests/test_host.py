"""
Host side of the boundary without a GPU: budget arithmetic of the tracers, parameter
blocks, pipeline addressing and error behaviour, checked against the reference's formulas
(src/theia/trace.py:623-634, 1176-1187; docs/pipeline/pipeline.md:57-64).
"""

import ctypes

import numpy as np
import pytest

import harness as H
from theia_b200 import _abi, _device, light, material, pipeline, random, response, target, trace


def build(builder, *a, **kw):
    with _device.useDevice("cpu"):
        return builder(*a, **kw)


@pytest.mark.parametrize("nScattering", [1, 6, 10])
@pytest.mark.parametrize("mis", [True, False])
@pytest.mark.parametrize("direct", [True, False])
@pytest.mark.parametrize("lightKind,nLight", [("spherical", 3), ("cone", 3), ("pencil", 1), ("cascade", 12), ("muon", 3)])
def test_volume_tracer_budgets(nScattering, mis, direct, lightKind, nLight):
    cfg = build(H.volumeScenario, 256, nScattering=nScattering, mis=mis, direct=direct, lightKind=lightKind,
                wavelength="uniform")
    tr = cfg["tracer"]
    hits = nScattering * (2 if mis else 1) + (1 if direct else 0)
    pathLength = nScattering - 1 if mis else nScattering
    assert tr.maxHitsPerThread == hits
    assert tr.maxHits == hits * 256
    assert tr.nRNGSamples == nLight + 1 + (7 if mis else 3) * pathLength
    assert tr.source.nRNGForward == nLight
    assert tr.batchSize == 256 and tr.capacity == 256 and tr.blockSize == 128
    assert tr.normalization == pytest.approx(1.0 / 256)
    assert cfg["response"].capacity == tr.maxHits


@pytest.mark.parametrize("guide", [False, True])
@pytest.mark.parametrize("direct", [True, False])
def test_scene_tracer_budgets(guide, direct):
    with _device.useDevice("cpu"):
        cfg = H.sceneC3(512, guide=guide, disableDirectLighting=not direct) if "guide" in H.sceneC3.__code__.co_varnames \
            else H.sceneC3(512, disableDirectLighting=not direct)
    tr = cfg["tracer"]
    hasGuide = tr.targetGuide is not None
    L = tr.maxPathLength
    hits = (L - 1) * (2 if hasGuide else 1) + (1 if direct else 0)
    assert tr.maxHitsPerThread == hits
    assert tr.nRNGSamples == 3 + 0 + (4 + (2 if hasGuide else 0)) * L


def test_batch_size_can_shrink_but_not_grow():
    cfg = build(H.volumeScenario, 1024)
    tr = cfg["tracer"]
    tr.batchSize = 100
    assert tr.batchSize == 100 and tr.capacity == 1024
    assert tr.normalization == pytest.approx(1.0 / 100)
    with pytest.raises(ValueError):
        tr.batchSize = 2048


def test_launch_description_of_a_volume_tracer():
    cfg = build(H.volumeScenario, 512, polarized=True, water=True, callback="stat")
    tr = cfg["tracer"]
    for _, st in tr.collectStages():
        st.update(0)
    L, keep = tr._launch(0)
    assert L.tracer == _abi.TRACER_VOLUME_FORWARD
    assert L.batch_size == 512 and L.polarized == 1
    assert L.light.kind == _abi.LIGHT_SPHERICAL and L.wavelength.kind == _abi.WAVELENGTH_CONST
    assert L.target.kind == _abi.TARGET_SPHERE
    assert L.response.kind == _abi.RESPONSE_RECORD and L.response.queue_capacity == tr.maxHits
    assert L.callback.kind == _abi.CALLBACK_STAT
    assert L.rng.key[0] == 42 and L.rng.key[1] == 0
    assert L.workspace is None  # volume tracer: register resident, no wavefront state


def test_philox_parameter_block_and_auto_advance():
    with _device.useDevice("cpu"):
        rng = random.PhiloxRNG(key=0x1122334455667788, offset=(1 << 96) + 5)
    rng.autoAdvance = 10
    rng.update(0)
    p = rng._block("PhiloxParams", 0) if hasattr(rng, "_block") else None
    raw = bytes(p)
    words = np.frombuffer(raw, np.uint32)
    assert words[0] == 0x55667788 and words[1] == 0x11223344
    # src/theia/random.py:278-282: the offset is advanced first, then uploaded
    assert list(words[2:6]) == [15, 0, 0, 1]
    rng.update(1)
    assert rng.getParam("offset") == (1 << 96) + 25


def test_pipeline_addresses_parameters_by_stage_name():
    cfg = build(H.volumeScenario, 256)
    tr = cfg["tracer"]
    pipe = pipeline.Pipeline(tr.collectStages())
    names = [n for n, _ in pipe.stages]
    assert names == ["rng", "photons", "lightSource", "target", "tracer", "response"][: len(names)] or "tracer" in names
    pipe.setParams(lightSource__budget=5.0, tracer__maxTime=123.0, target__radius=0.25)
    assert tr.source.getParam("budget") == 5.0
    assert tr.getParam("maxTime") == 123.0
    assert tr.target.getParam("radius") == 0.25
    got = pipe.getParams()
    assert got["lightSource__budget"] == 5.0 and got["tracer__maxTime"] == 123.0


def test_tracers_refuse_to_run_without_a_gpu():
    cfg = build(H.volumeScenario, 256)
    with pytest.raises(RuntimeError):
        pipeline.runPipeline(cfg["tracer"].collectStages())
    with pytest.raises(RuntimeError):
        with _device.useDevice("cuda"):
            pass


def test_material_flags():
    # src/theia/material.py:531-557
    F = material.MaterialFlags
    P = material.parseMaterialFlags
    assert P("") == F.NO_REFLECT | F.NO_TRANSMIT
    assert P("T") == F.NO_REFLECT and P("R") == F.NO_TRANSMIT and P("TR") == F(0)
    assert P("Tf") == F.NO_REFLECT | F.NO_TRANSMIT_BWD
    assert P("DB") == F.DETECTOR | F.BLACK_BODY | F.NO_REFLECT | F.NO_TRANSMIT
    assert P("V") & F.VOLUME_BORDER
    with pytest.raises(ValueError):
        material.parseMaterialFlags("X")
    m = material.Material("m", None, None, flags=("TR", "B"))
    assert m.flagsInward == material.parseMaterialFlags("TR") and m.flagsOutward == material.parseMaterialFlags("B")


def test_hit_queue_layout():
    # sh/response.queue.glsl:6-33: count, then field-major arrays of length capacity
    from theia_b200.queue import QueueView
    n = 5
    raw = np.zeros(4 + 13 * 4 * n, np.uint8)
    v = QueueView(raw, response.HitItem, n)
    assert v.fields[:3] == ["position", "direction", "normal"]
    v.count = 3
    v["position"][1] = (1.0, 2.0, 3.0)
    f = raw[4:].view(np.float32)
    assert raw[:4].view(np.uint32)[0] == 3
    assert f[1] == 1.0 and f[n + 1] == 2.0 and f[2 * n + 1] == 3.0
    assert v[1:3]["position"].shape == (2, 3) and v[1:3].count == 2


def test_scheduler_with_named_pipelines():
    # examples/03_multiple_lightsources.ipynb cell 6-8: tasks are (pipeline name, params)
    class Stage(pipeline.PipelineStage):
        name = "stage"

        class P(ctypes.Structure):
            _fields_ = [("value", ctypes.c_float)]

        def __init__(self, log, tag):
            super().__init__({"P": self.P})
            self._log, self._tag = log, tag

        def run(self, i):
            return [lambda: self._log.append((self._tag, i, self._block("P", i).value))]

    log = []
    with _device.useDevice("cpu"):
        a, b = Stage(log, "a"), Stage(log, "b")
        pa, pb = pipeline.Pipeline([("s", a)]), pipeline.Pipeline([("s", b)])
    done = []
    sched = pipeline.PipelineScheduler({"a": pa, "b": pb}, processFn=lambda i, n, args: done.append((i, n)))
    sched.schedule([("a", {"s__value": 1.0}), ("b", {"s__value": 2.0}), ("a", {}), ("b", {"s__value": 4.0})])
    sched.wait()
    assert log == [("a", 0, 1.0), ("b", 1, 2.0), ("a", 0, 1.0), ("b", 1, 4.0)]
    assert done == [(0, 0), (1, 1), (0, 2), (1, 3)]
    with pytest.raises(ValueError):
        sched.schedule([{"s__value": 1.0}])
    with pytest.raises(KeyError):
        sched.schedule([("c", {})])
    sched.destroy()


def test_converge_histogram_task_on_a_stub_pipeline():
    # src/theia/task.py:22-196: batches are issued until the standard error of the bin sum is small
    from theia_b200 import task

    class FakeResponse:
        nBins, retrieve = 4, True

        def __init__(self):
            self.rng = np.random.default_rng(5)
            self.buffers = [np.zeros(4), np.zeros(4)]

        def result(self, i):
            return self.buffers[i]

    resp = FakeResponse()

    class Producer(pipeline.PipelineStage):
        name = "producer"

        class P(ctypes.Structure):
            _fields_ = [("sigma", ctypes.c_float)]

        def __init__(self):
            super().__init__({"P": self.P})

        def run(self, i):
            def cmd():
                resp.buffers[i][:] = 10.0 + self._block("P", i).sigma * resp.rng.standard_normal(4)
            return [cmd]

    with _device.useDevice("cpu"):
        pipe = pipeline.Pipeline([("src", Producer())])
    finished = []
    sched = pipeline.DynamicTaskScheduler(pipe)
    quiet = task.ConvergeHistogramTask(resp, {"src__sigma": 0.01}, atol=0.05, rtol=0.0, finishedCallback=finished.append)
    noisy = task.ConvergeHistogramTask(resp, {"src__sigma": 2.0}, atol=0.5, rtol=0.0, maxBatchCount=200,
                                       finishedCallback=finished.append)
    sched.scheduleTasks([quiet, noisy])
    sched.wait()
    sched.destroy()
    assert quiet.converged and quiet.totalBatches == 4 and quiet.isFinished
    assert noisy.converged and noisy.totalBatches > 4 and noisy.error <= 0.5
    assert abs(noisy.result.sum() - 40.0) < 3 * 0.5
    assert finished == [quiet, noisy] or finished == [noisy, quiet]
    with pytest.raises(ValueError):
        task.ConvergeHistogramTask(resp, {}, initialBatchCount=1)
    hopeless = task.ConvergeHistogramTask(resp, {"src__sigma": 5.0}, atol=1e-6, rtol=0.0, maxBatchCount=6)
    sched = pipeline.DynamicTaskScheduler(pipe)
    with pytest.warns(UserWarning, match="Failed to converge"):
        sched.scheduleTask(hopeless)
        sched.wait()
    sched.destroy()
    assert not hopeless.converged and hopeless.totalBatches == 6


def test_host_sources_stream_batches():
    # src/theia/light.py:147-255, 775-880: Streaming*Source walk through the data batch by batch;
    # new data rewinds; running dry warns and leaves the queue as it is
    import warnings
    from theia_b200 import light
    with _device.useDevice("cpu"):
        rays, photons = H.hostSourceData(96, polarized=True)
        src = light.StreamingHostLightSource(64, data=rays, polarized=True)
        wl = light.StreamingHostWavelengthSource(64, data=photons, ignoreContrib=True)
        assert src.queue.view(0).fields == ["position", "direction", "stokes", "polarizationRef", "startTime", "contrib"]
        assert src.queue.buffer(0).numel() == 4 * 15 * 64  # no count word (skipCounter)
        src.update(0); wl.update(0)
        assert np.array_equal(src.queue.view(0)["direction"], rays["direction"][:64])
        assert np.array_equal(wl.queue.view(0)["wavelength"], photons["wavelength"][:64])
        assert np.all(wl.queue.view(0)["contrib"] == 1.0)
        src.update(1)
        assert np.array_equal(src.queue.view(1)["startTime"][:32], rays["startTime"][64:])
        with warnings.catch_warnings(record=True) as w:
            warnings.simplefilter("always")
            src.update(0)
        assert any("ran out of data" in str(x.message) for x in w)
        src.data = rays  # rewind
        src.update(1)
        assert np.array_equal(src.queue.view(1)["position"], rays["position"][:64])
        with pytest.raises(ValueError):
            src.data = {"position": rays["position"]}
        # the parameter block handed to the C-ABI names the device queue
        blk = src._firstBlock(1)
        assert blk._queue == src.queue.tensor.data_ptr() and blk._capacity == 64 and blk._polarized == 1
        for cmd in src.run(1):
            cmd()
        dev = src.queue.tensor.numpy().view(np.float32).reshape(15, 64)
        assert np.array_equal(dev[3:6].T, rays["direction"][:64])


def test_host_sources_on_the_oracle():
    # a path starts exactly where the host says: first track vertex = queue entry
    with _device.useDevice("cpu"):
        cfg = H.volumeScenario(256, lightKind="host", wavelength="host", callback="track", polarized=True, water=True)
        H.runOracle(cfg)
        q = cfg["tracer"].source.queue.view(0)
        tracks, lengths, codes = cfg["callback"].result(0)
        assert np.array_equal(tracks[:, 0, :3], q["position"])
        assert np.array_equal(tracks[:, 0, 3], q["startTime"])
        hits = H.hitView(cfg)
        wl = cfg["tracer"].wavelengthSource.queue.view(0)["wavelength"]
        assert hits.count > 10 and np.all(np.isin(hits["wavelength"], wl))


# ------------------------------------------------------------------ tests/test_lookup.py of the reference (host side)
def _sampleData1D():
    x = np.linspace(0.0, 1.0, 100)
    y = np.sin(4 * np.pi * x) * np.exp(-x) + (1.0 - x)
    grad = -np.exp(-x) * (np.exp(x) + np.sin(4 * np.pi * x) - 4 * np.pi * np.cos(4 * np.pi * x))
    return np.stack((x, y, grad), axis=-1)


def _sampleData2D():
    x, y = np.meshgrid(np.linspace(0.0, 1.0, 50), np.linspace(0.0, 1.0, 100))
    x, y = x.flatten(), y.flatten()
    z = np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y) * (1.0 - y * y) * (1.0 - x * x)
    return np.stack((x, y, z), axis=-1)


def test_sampleTable1D():  # tests/test_lookup.py:20-36
    from theia_b200 import lookup
    data = np.array([[0.0, 0.0], [10.0, 10.0]])
    sample = lookup.sampleTable1D(data, 10).data
    assert abs(sample.min()) < 1e-5 and abs(sample.max() - 10.0) < 1e-5 and len(sample) == 10
    sample = lookup.sampleTable1D(data, 10, boundary=(5.0, 8.0)).data
    assert abs(sample.min() - 5.0) < 1e-5 and abs(sample.max() - 8.0) < 1e-5 and len(sample) == 10
    sample = lookup.sampleTable1D(_sampleData1D(), 1024).data
    x = np.linspace(0.0, 1.0, 1024)
    y = np.sin(4 * np.pi * x) * np.exp(-x) + (1.0 - x)
    assert np.abs(y - sample).max() < 5e-3


def test_sampleTable2D():  # tests/test_lookup.py:117-138
    from theia_b200 import lookup
    data_x = np.array([0.0, 0.0, 10.0, 10.0])
    data_y = np.array([0.0, 10.0, 0.0, 10.0])
    data = np.stack((data_x, data_y, data_x + data_y), axis=-1)
    sample = lookup.sampleTable2D(data, 100, 100, boundaries=(None, (3.0, 8.0))).data.flatten()
    assert sample.size == 100 * 100
    assert abs(sample[:100].min() - 3) < 1e-5 and abs(sample[:100].max() - 13.0) < 1e-5
    assert abs(sample[::100].min() - 3.0) < 1e-5 and abs(sample[::100].max() - 8.0) < 1e-5
    sample = lookup.sampleTable2D(_sampleData2D(), 250, 250, mode="cubic").data.flatten()
    x, y = np.meshgrid(np.linspace(0.0, 1.0, 250), np.linspace(0.0, 1.0, 250))
    z = np.sin(2 * np.pi * x) * np.cos(2 * np.pi * y) * (1.0 - y * y) * (1.0 - x * x)
    assert np.abs(z.flatten() - sample).max() < 5e-3


def test_getTableSize():  # tests/test_lookup.py:176-184
    from theia_b200 import lookup
    for column in (_sampleData1D()[:, 1], _sampleData2D()[:, 2]):
        assert lookup.getTableSize(column) == lookup.Table(column).nbytes


def test_uploadTables():  # tests/test_lookup.py:187-214
    from theia_b200 import _device, lookup
    rng = np.random.default_rng(0xC0FFEE)
    data = [rng.random((64,)), rng.random((16,)), rng.random((32,))]
    with _device.useDevice("cpu"):
        tensor, ptr = lookup.uploadTables(data)
    tables = [lookup.Table(t) for t in data]
    expected_ptr = tensor.data_ptr()
    for t, p in zip(tables, ptr):
        assert p == expected_ptr
        expected_ptr += t.nbytes
    size = sum(t.nbytes for t in tables)
    expected = np.empty(size, dtype=np.uint8)
    at = expected.ctypes.data
    for t in tables:
        at += t.copy(at)
    assert np.array_equal(tensor.numpy()[:size], expected)


def test_tracerReportsConfig():  # tests/test_trace.py:22-79
    from theia_b200 import _device, light, random, response, target, trace
    batchSize, batchSize2, capacity, polarized = 64 * 1024, 96 * 1024, 128 * 1024, True

    class MockResponse(response.EmptyResponse):
        def __init__(self) -> None:
            super().__init__()
            self.prepareConfig = None
            self.updatedConfig = None

        def prepare(self, config):
            self.prepareConfig = config

        def updateConfig(self, config):
            self.updatedConfig = config

    with _device.useDevice("cpu"):
        resp = MockResponse()
        tracer = trace.VolumeForwardTracer(batchSize, light.SphericalLightSource(), target.SphereTarget(),
                                           light.ConstWavelengthSource(), resp, rng=random.PhiloxRNG(key=0xABBA), medium=0,
                                           capacity=capacity, polarized=polarized)
        c = resp.prepareConfig
        assert c is not None
        assert (c.batchSize, c.blockSize, c.capacity) == (batchSize, tracer.blockSize, capacity)
        assert c.maxHitsPerThread == tracer.maxHitsPerThread
        assert c.normalization == tracer.normalization == 1.0 / batchSize
        assert c.polarized == polarized
        tracer.batchSize = batchSize2
        tracer.update(0)
        c = resp.updatedConfig
        assert c is not None
        assert (c.batchSize, c.blockSize, c.capacity) == (batchSize2, tracer.blockSize, capacity)
        assert c.maxHitsPerThread == tracer.maxHitsPerThread
        assert c.normalization == tracer.normalization == 1.0 / batchSize2
        assert c.polarized == polarized
