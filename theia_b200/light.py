"""
Light and wavelength sources (API of the reference's src/theia/light.py, forward path).

Each class owns a parameter block that is byte-identical to the reference's
``ctypes.Structure`` and to the matching ``theia_light_*`` / ``theia_wavelength_*`` struct of
include/theia_b200.h; the device code is selected by ``_kind`` / ``_flags``.
"""

from __future__ import annotations

from ctypes import Structure, c_float, c_uint32, c_uint64
from typing import Callable
from warnings import warn

import numpy as np

import theia_b200.units as u
from theia_b200 import _abi
from theia_b200.glsl import buffer_reference, vec2, vec3, vec4
from theia_b200.pipeline import PipelineStage, SourceCodeMixin

__all__ = [
    "CherenkovLightSource", "CherenkovTrackLightSource", "ConeLightSource", "ConstWavelengthSource",
    "FunctionWavelengthSource", "HostLightSource", "HostWavelengthSource", "LightSampleItem", "ParticleTrack",
    "LightSampler", "LightSource", "MuonTrackLightSource", "ParticleCascadeLightSource", "PencilLightSource",
    "PolarizedLightSampleItem", "SphericalLightSource", "StreamingHostLightSource",
    "StreamingHostWavelengthSource", "UniformWavelengthSource", "WavelengthSampleItem", "WavelengthSource",
    "frankTamm",
]


# ------------------------------------------------------------------ wavelength sources
class WavelengthSource(SourceCodeMixin):
    """Base class of wavelength samplers (src/theia/light.py:59-83)."""

    name = "Wavelength Sampler"

    def __init__(self, *, nRNGSamples: int = 0, params: dict[str, type[Structure]] = {},
                 extra: set[str] = set()) -> None:
        super().__init__(params, extra)
        self._nRNGSamples = nRNGSamples

    @property
    def nRNGSamples(self) -> int:
        """Amount of random numbers drawn per sample"""
        return self._nRNGSamples


class WavelengthSampleItem(Structure):
    _fields_ = [("wavelength", c_float), ("contrib", c_float)]


class _HostQueueParams(Structure):
    # theia_light_host_t / theia_wavelength_host_t of include/theia_b200.h (the reference binds the
    # queue tensor by name instead: program.bindParams(LightQueueIn=...), src/theia/light.py:761)
    _pack_ = 4
    _fields_ = [("_queue", c_uint64), ("_capacity", c_uint32), ("_polarized", c_uint32)]


def _streamInto(source, i: int, what: str) -> None:
    """next ``capacity`` entries of source.data -> host copy i of the queue (src/theia/light.py:232-255, 856-880)"""
    if source.data is None:
        warn(f"{what}: New data was requested but no source was given!")
        return
    if source._offset >= source._dataSize:
        warn(f"{what} ran out of data!")
        return
    start = source._offset
    end = min(source._offset + source.capacity, source._dataSize)
    source._offset = end
    view = source.queue.view(i)
    for field in source._dataFields:
        view[field][: end - start] = source.data[field][start:end]


class HostWavelengthSource(WavelengthSource):
    """
    Wavelength source passing samples from the host to the device: path k of a batch reads
    entry k of ``queue`` (src/theia/light.py:87-144, sh/wavelengthsource.host.glsl). ``updateFn(source, i)``
    is called before configuration i is committed and may refill ``queue.view(i)``.
    """

    name = "Host Wavelength Source"
    _kind = _abi.WAVELENGTH_HOST

    def __init__(self, capacity: int, *, updateFn: Callable[["HostWavelengthSource", int], None] | None = None,
                 extra: set[str] = set()) -> None:
        from theia_b200.queue import IOQueue
        super().__init__(params={"HostParams": _HostQueueParams}, extra=extra)
        self._capacity = capacity
        self._updateFn = updateFn
        self._queue = IOQueue(WavelengthSampleItem, capacity, mode="update", skipCounter=True)
        self.setParams(_queue=self._queue.tensor.data_ptr(), _capacity=capacity, _polarized=0)

    @property
    def capacity(self) -> int:
        """Maximum number of samples that can be drawn per run"""
        return self._capacity

    @property
    def queue(self):
        """Queue containing the data for the next batch"""
        return self._queue

    def update(self, i: int) -> None:
        if self._updateFn is not None:
            self._updateFn(self, i)
        super().update(i)

    def run(self, i: int):
        return [*self.queue.run(i), *super().run(i)]


class StreamingHostWavelengthSource(HostWavelengthSource):
    """
    Feeds pre-sampled wavelengths batch by batch from ``data`` (dict with ``wavelength`` and,
    unless ``ignoreContrib``, ``contrib``); assigning new data rewinds the stream
    (src/theia/light.py:147-255). Entries beyond the end of the data keep their old content.
    """

    name = "Streaming Host Wavelength Source"

    def __init__(self, batchSize: int, *, data=None, ignoreContrib: bool = False) -> None:
        super().__init__(batchSize, updateFn=self._updateData, extra={"data"})
        if ignoreContrib:
            self._dataFields = {"wavelength"}
            self.queue.view(0)["contrib"] = 1.0
            self.queue.view(1)["contrib"] = 1.0
        else:
            self._dataFields = {"contrib", "wavelength"}
        self.data = data

    @property
    def data(self):
        """Data used as streaming source"""
        return self._data

    @data.setter
    def data(self, value) -> None:
        self._offset = 0
        self._dataSize = 0
        if value is None:
            self._data = None
            return
        if not self._dataFields.issubset(value.keys()):
            raise ValueError(f"provided data misses expected fields: {self._dataFields.difference(value.keys())}")
        self._dataSize = np.size(value["wavelength"], 0)
        for field in self._dataFields:
            if np.size(value[field], 0) != self._dataSize:
                raise ValueError("fields of provided data differ in size!")
        self._data = {f: np.asarray(value[f]) for f in self._dataFields}

    @data.deleter
    def data(self) -> None:
        self.data = None

    @staticmethod
    def _updateData(source: HostWavelengthSource, i: int) -> None:
        _streamInto(source, i, "StreamingHostWavelengthSource")


class FunctionWavelengthSource(WavelengthSource):
    """
    Importance samples wavelengths from ``fn`` (nm -> density): the distribution is integrated
    and inverted numerically on the host and uploaded as a look-up table of the inverse CDF
    (src/theia/light.py:351-414, sh/wavelengthsource.function.glsl). The contribution of every
    sample is the integral of ``fn`` over ``lambdaRange``.
    """

    name = "Function Wavelength Source"
    _kind = _abi.WAVELENGTH_FUNCTION

    class WavelengthParams(Structure):
        _pack_ = 4
        _fields_ = [("_table", c_uint64), ("_contrib", c_float)]

    def __init__(self, fn: Callable[[float], float], *, lambdaRange: tuple[float, float] = (300.0, 700.0),
                 numSamples: int = 1024) -> None:
        super().__init__(nRNGSamples=1, params={"WavelengthParams": self.WavelengthParams})
        self._updateFn(fn, lambdaRange, numSamples)

    def _updateFn(self, fn, lambdaRange, numSamples: int) -> None:
        from scipy.integrate import quad
        from scipy.stats.sampling import NumericalInversePolynomial
        from theia_b200.lookup import uploadTables

        contrib, _ = quad(fn, *lambdaRange)

        class Dist:
            def pdf(self, x):
                return fn(x)

        inv_cdf = NumericalInversePolynomial(Dist(), domain=lambdaRange)
        x = inv_cdf.ppf(np.linspace(0.0, 1.0, numSamples))
        self._tableMemory, [adr] = uploadTables([x])
        self.setParams(_table=adr, _contrib=contrib)


class ConstWavelengthSource(WavelengthSource):
    """Constant wavelength (src/theia/light.py:258-283, sh/wavelengthsource.const.glsl)."""

    name = "Const Wavelength Source"
    _kind = _abi.WAVELENGTH_CONST

    class WavelengthParams(Structure):
        _fields_ = [("wavelength", c_float)]

    def __init__(self, wavelength: float = 600.0 * u.nm) -> None:
        super().__init__(params={"WavelengthParams": self.WavelengthParams})
        self.setParams(wavelength=wavelength)


class UniformWavelengthSource(WavelengthSource):
    """
    Uniform wavelengths on ``lambdaRange`` (src/theia/light.py:286-348). With
    ``normalize=False`` the contribution carries the width of the range.
    """

    name = "Uniform Wavelength Source"
    _kind = _abi.WAVELENGTH_UNIFORM

    class SourceParams(Structure):
        _fields_ = [("lambdaRange", vec2), ("_contrib", c_float)]

    def __init__(self, *, lambdaRange: tuple[float, float] = (300.0, 700.0), normalize: bool = True) -> None:
        super().__init__(nRNGSamples=1, params={"WavelengthParams": self.SourceParams}, extra={"normalize"})
        self.setParams(lambdaRange=lambdaRange, normalize=normalize)

    @property
    def normalize(self) -> bool:
        return self._normalize

    @normalize.setter
    def normalize(self, value: bool) -> None:
        self._normalize = value

    def _finishParams(self, i: int) -> None:
        lo, hi = self.getParam("lambdaRange")
        width = hi - lo
        c = 1.0
        if width != 0.0 and not self.normalize:
            c *= abs(width)
        self.setParam("_contrib", c)


# ------------------------------------------------------------------ light sources
class LightSource(SourceCodeMixin):
    """Base class of light sources (src/theia/light.py:416-460)."""

    name = "Light Source"

    def __init__(self, *, supportForward: bool, supportBackward: bool, nRNGForward: int = 0, nRNGBackward: int = 0,
                 params: dict[str, type[Structure]] = {}, extra: set[str] = set()) -> None:
        super().__init__(params, extra)
        self._forward = supportForward
        self._backward = supportBackward
        self._rngForward = nRNGForward
        self._rngBackward = nRNGBackward

    @property
    def supportForward(self) -> bool:
        return self._forward

    @property
    def supportBackward(self) -> bool:
        """Backward sampling belongs to the camera side tracers, which this build omits."""
        return self._backward

    @property
    def nRNGForward(self) -> int:
        return self._rngForward

    @property
    def nRNGBackward(self) -> int:
        return self._rngBackward


class LightSampleItem(Structure):
    _fields_ = [("position", c_float * 3), ("direction", c_float * 3), ("startTime", c_float), ("contrib", c_float)]


class PolarizedLightSampleItem(Structure):
    _fields_ = [("position", c_float * 3), ("direction", c_float * 3), ("stokes", c_float * 4),
                ("polarizationRef", c_float * 3), ("startTime", c_float), ("contrib", c_float)]


class LightSampler(PipelineStage):
    """
    Samples a ``LightSource`` (+ ``WavelengthSource``) into queues (src/theia/light.py:492-690,
    sh/lightsource.sample.glsl): sample k of a run is stream ``baseCount + k`` of the generator, drawing
    the wavelength first and the light ray second, exactly like path k of a tracer does.

    Parameters: ``source, wavelengthSource, capacity``; keywords ``rng, medium, polarized, retrieve,
    scene, batchSize``. Stage parameters: ``count`` (samples per run, at most ``capacity``),
    ``baseCount`` (offset into the sampling stream), ``medium`` (address of the medium the source is
    submerged in; vacuum if 0).
    """

    name = "Light Sampler"

    class SampleParams(Structure):
        _fields_ = [("count", c_uint32), ("baseCount", c_uint32), ("medium", buffer_reference)]

    def __init__(self, source: LightSource, wavelengthSource: WavelengthSource, capacity: int, *, rng=None,
                 medium: int | None = None, polarized: bool = False, retrieve: bool = True, scene=None,
                 batchSize: int = 128, code: bytes | None = None) -> None:
        from theia_b200.queue import IOQueue
        needRNG = source.nRNGForward > 0 or wavelengthSource.nRNGSamples > 0
        if needRNG and rng is None:
            raise ValueError("Light source requires a RNG but none was given!")
        if isinstance(source, HostLightSource) and source.polarized != polarized:
            raise RuntimeError("mismatch in light queue definition")  # sh/lightsource.sample.glsl:25-29
        super().__init__({"SampleParams": self.SampleParams})
        if medium is None:
            medium = 0 if scene is None else scene.medium
        self._batchSize = batchSize
        self._capacity = capacity
        self._scene = scene
        self._source = source
        self._retrieve = retrieve
        self._polarized = polarized
        self._rng = rng
        self._wavelengthSource = wavelengthSource
        self._code = code or b""
        self.setParams(count=capacity, baseCount=0, medium=medium)
        self._lightQueue = IOQueue(PolarizedLightSampleItem if polarized else LightSampleItem, capacity, mode="retrieve")
        self._lamQueue = IOQueue(WavelengthSampleItem, capacity, mode="retrieve")

    @property
    def batchSize(self) -> int:
        """Number of samples drawn per work group"""
        return self._batchSize

    @property
    def capacity(self) -> int:
        """Maximum number of samples that can be drawn per run"""
        return self._capacity

    @property
    def code(self) -> bytes:
        return self._code

    @property
    def lightQueue(self):
        """Queue storing the light samples"""
        return self._lightQueue

    @property
    def polarized(self) -> bool:
        return self._polarized

    @property
    def retrieve(self) -> bool:
        return self._retrieve

    @property
    def rng(self):
        return self._rng

    @property
    def scene(self):
        return self._scene

    @property
    def source(self) -> LightSource:
        return self._source

    @property
    def wavelengthSource(self) -> WavelengthSource:
        return self._wavelengthSource

    @property
    def wavelengthQueue(self):
        """Queue holding the sampled wavelengths"""
        return self._lamQueue

    def collectStages(self) -> list[PipelineStage]:
        stages = [] if self.rng is None else [self.rng]
        stages.extend([self.wavelengthSource, self.source, self])
        return stages

    def _unitArgs(self, i: int):
        """Launch description of configuration i (theia_unit_eval, unit LIGHT_SAMPLE)"""
        from theia_b200 import _units
        p = self._block("SampleParams", i)
        if p.count > self._capacity:
            raise ValueError("count exceeds the capacity of the sampler")
        A = _units.unitArgs(_abi.UNIT_LIGHT_SAMPLE, p.count, rng=self._rng, i=i, polarized=self._polarized,
                            first=p.baseCount, medium=p.medium.value, light=self._source, wavelength=self._wavelengthSource)
        A.light_queue = self._lightQueue.tensor.data_ptr()
        A.wavelength_queue = self._lamQueue.tensor.data_ptr()
        A.capacity = self._capacity
        return A

    def run(self, i: int):
        from ctypes import byref
        from theia_b200 import _device, _lib

        def cmd() -> None:
            if self._device != "cuda":
                raise RuntimeError("LightSampler needs a CUDA device (no CPU fallback)")
            A = self._unitArgs(i)
            A.stream = _device.currentStream("cuda")
            _lib.check(_lib.load().theia_unit_eval(byref(A)))

        cmds = [cmd]
        if self._retrieve:
            cmds += [*self._lightQueue.run(i), *self._lamQueue.run(i)]
        return cmds


class HostLightSource(LightSource):
    """
    Light source passing rays from the host to the device: path k of a batch starts from entry k
    of ``queue`` (src/theia/light.py:692-772, sh/lightsource.host.glsl). Forward mode only.
    ``updateFn(source, i)`` is called before configuration i is committed.
    """

    name = "Host Light Source"
    _kind = _abi.LIGHT_HOST

    def __init__(self, capacity: int, *, polarized: bool = False,
                 updateFn: Callable[["HostLightSource", int], None] | None = None, extra: set[str] = set()) -> None:
        from theia_b200.queue import IOQueue
        super().__init__(supportForward=True, supportBackward=False, params={"HostParams": _HostQueueParams}, extra=extra)
        self._capacity = capacity
        self._polarized = polarized
        self._updateFn = updateFn
        item = PolarizedLightSampleItem if polarized else LightSampleItem
        self._queue = IOQueue(item, capacity, mode="update", skipCounter=True)
        self.setParams(_queue=self._queue.tensor.data_ptr(), _capacity=capacity, _polarized=int(polarized))

    @property
    def capacity(self) -> int:
        """Maximum number of samples that can be drawn per run"""
        return self._capacity

    @property
    def polarized(self) -> bool:
        """Whether the light source provides polarization information"""
        return self._polarized

    @property
    def queue(self):
        """Queue containing the data for the next batch"""
        return self._queue

    @property
    def _flags(self) -> int:
        return _abi.LIGHT_FLAG_POLARIZED if self._polarized else 0

    def update(self, i: int) -> None:
        if self._updateFn is not None:
            self._updateFn(self, i)
        super().update(i)

    def run(self, i: int):
        return [*self.queue.run(i), *super().run(i)]


class StreamingHostLightSource(HostLightSource):
    """
    Feeds light rays batch by batch from ``data`` (dict of arrays named like the fields of
    ``LightSampleItem`` / ``PolarizedLightSampleItem``); assigning new data rewinds the stream
    (src/theia/light.py:775-880).
    """

    name = "Streaming Host Light Source"

    def __init__(self, batchSize: int, *, data=None, polarized: bool = False, ignoreContrib: bool = False) -> None:
        super().__init__(batchSize, polarized=polarized, updateFn=self._updateData, extra={"data"})
        self._dataFields = set(self.queue.view(0).fields)
        if ignoreContrib:
            self._dataFields.remove("contrib")
            self.queue.view(0)["contrib"] = 1.0
            self.queue.view(1)["contrib"] = 1.0
        self.data = data

    @property
    def data(self):
        """Data used as streaming source"""
        return self._data

    @data.setter
    def data(self, value) -> None:
        self._offset = 0
        self._dataSize = 0
        if value is None:
            self._data = None
            return
        if not self._dataFields.issubset(value.keys()):
            raise ValueError(f"provided data misses expected fields: {self._dataFields.difference(value.keys())}")
        self._dataSize = np.size(value["startTime"], 0)
        for field in self._dataFields:
            if np.size(value[field], 0) != self._dataSize:
                raise ValueError("fields of provided data differ in size!")
        self._data = {f: np.asarray(value[f]) for f in self._dataFields}

    @data.deleter
    def data(self) -> None:
        self.data = None

    @staticmethod
    def _updateData(source: HostLightSource, i: int) -> None:
        _streamInto(source, i, "StreamingHostLightSource")


class CherenkovLightSource(LightSource):
    """
    Cherenkov light of a straight track between ``trackStart`` and ``trackEnd`` travelled at the
    speed of light (beta = 1): emission points uniform along the track, directions on the
    Cherenkov cone of the medium's refractive index at the sampled wavelength; polarised tracers
    get linearly polarised light. The contribution is the Frank-Tamm yield in eV or, with
    ``usePhotonCount``, in photons (src/theia/light.py:1183-1271,
    sh/lightsource.cherenkov.simple.glsl). Forward sampling only in this build.
    """

    name = "Cherenkov Light Source"
    _kind = _abi.LIGHT_CHERENKOV

    class LightParams(Structure):
        _fields_ = [("trackStart", vec3), ("startTime", c_float), ("trackEnd", vec3), ("endTime", c_float),
                    ("_trackDir", vec3), ("_trackDist", c_float)]

    def __init__(self, *, trackStart=(0.0, 0.0, 0.0), trackEnd=(100.0, 0.0, 0.0), startTime: float = 0.0 * u.ns,
                 endTime: float = 100.0 * u.m / u.c, usePhotonCount: bool = False) -> None:
        super().__init__(supportForward=True, supportBackward=True, nRNGForward=2, nRNGBackward=0,
                         params={"LightParams": self.LightParams})
        self._usePhotonCount = usePhotonCount
        self.setParams(trackStart=trackStart, trackEnd=trackEnd, startTime=startTime, endTime=endTime)

    @property
    def usePhotonCount(self) -> bool:
        """Whether to sample radiance in eV or #photons"""
        return self._usePhotonCount

    @property
    def _flags(self) -> int:
        return _abi.LIGHT_FLAG_PHOTON_COUNT if self._usePhotonCount else 0

    def _finishParams(self, i: int) -> None:
        d = np.array(self.getParam("trackEnd"), dtype=np.float64) - np.array(self.getParam("trackStart"), dtype=np.float64)
        dist = np.sqrt(np.square(d).sum(-1))
        self.setParams(_trackDir=d / dist, _trackDist=dist)


class ParticleTrack:
    """
    Particle track on the device: ``uint32 length`` (number of segments) followed by ``capacity``
    vertices ``(x, y, z, t)`` (src/theia/light.py:1274-1353). The reference maps one tensor into
    host and device address space; here a host copy is edited and ``setVertices`` / ``flush``
    push it to the device.
    """

    class Header(Structure):
        _fields_ = [("length", c_uint32)]

    class Vertex(Structure):
        _fields_ = [("x", c_float), ("y", c_float), ("z", c_float), ("t", c_float)]

    def __init__(self, capacity: int) -> None:
        from theia_b200 import _device
        self._capacity = capacity
        self._host = np.zeros(1 + 4 * capacity, dtype=np.uint32)
        self._tensor = _device.alloc(self._host.nbytes, _device.currentDevice())

    @property
    def capacity(self) -> int:
        """Maximum number of vertices this track can store"""
        return self._capacity

    @property
    def address(self) -> int:
        """Device address of the track"""
        return self._tensor.data_ptr()

    @property
    def length(self) -> int:
        """Number of segments in the track, i.e. number of vertices - 1"""
        return int(self._host[0])

    @length.setter
    def length(self, value: int) -> None:
        if not 0 <= value <= self.capacity - 1:
            raise ValueError("length must be between 0 and (capacity - 1)")
        self._host[0] = value

    def numpy(self, flat: bool = False) -> np.ndarray:
        """Host copy of the vertices: structured, or ``(capacity, 4)`` floats if ``flat``"""
        f = self._host[1:].view(np.float32).reshape(-1, 4)
        return f if flat else f.view([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("t", "<f4")]).reshape(-1)

    def flush(self) -> None:
        """Copies the host copy to the device"""
        import torch
        self._tensor[: self._host.nbytes].copy_(torch.from_numpy(self._host.view(np.uint8)))

    def setVertices(self, data) -> None:
        """Copies vertices of shape (length + 1, 4) = (x[m], y[m], z[m], t[ns]) into the track"""
        data = np.asarray(data, dtype=np.float32)
        self.numpy(True)[: len(data)] = data
        self.length = len(data) - 1
        self.flush()


class CherenkovTrackLightSource(LightSource):
    """
    Cherenkov light of a piecewise linear ``ParticleTrack`` (beta = 1): a segment is picked
    uniformly, then sampled like ``CherenkovLightSource``
    (src/theia/light.py:1356-1410, sh/lightsource.cherenkov.track.glsl).
    """

    name = "Cherenkov Track Light Source"
    _kind = _abi.LIGHT_CHERENKOV_TRACK

    class TrackParams(Structure):
        _fields_ = [("track", buffer_reference)]

    def __init__(self, track: "ParticleTrack | int | None" = None, *, usePhotonCount: bool = False) -> None:
        super().__init__(supportForward=True, supportBackward=False, nRNGForward=2,
                         params={"TrackParams": self.TrackParams})
        self._usePhotonCount = usePhotonCount
        self._trackObject = track if isinstance(track, ParticleTrack) else None
        adr = track.address if isinstance(track, ParticleTrack) else (track if track is not None else 0)
        self.setParams(track=adr)

    @property
    def usePhotonCount(self) -> bool:
        return self._usePhotonCount

    @property
    def _flags(self) -> int:
        return _abi.LIGHT_FLAG_PHOTON_COUNT if self._usePhotonCount else 0


class ConeLightSource(LightSource):
    """
    Point source radiating uniformly into a cone, optionally with a fixed polarisation
    (src/theia/light.py:883-1021, sh/lightsource.cone.glsl).
    """

    name = "Cone Light Source"
    _kind = _abi.LIGHT_CONE

    class LightParams(Structure):
        _fields_ = [
            ("position", vec3), ("direction", vec3), ("cosOpeningAngle", c_float), ("_contribFwd", c_float),
            ("_contribBwd", c_float), ("timeRange", vec2), ("polarizationReference", vec3), ("stokes", vec4),
        ]

    def __init__(self, *, position=(0.0, 0.0, 0.0), direction=(1.0, 0.0, 0.0), timeRange=(0.0, 100.0) * u.ns,
                 cosOpeningAngle: float = 0.5, budget: float = 1.0, polarizationReference=(0.0, 1.0, 0.0),
                 stokes=(1.0, 0.0, 0.0, 0.0), polarized: bool = True) -> None:
        super().__init__(supportForward=True, supportBackward=True, nRNGForward=3, nRNGBackward=1,
                         params={"LightParams": ConeLightSource.LightParams}, extra={"budget"})
        self._polarized = polarized
        self.budget = budget
        self.setParams(position=position, direction=direction, timeRange=timeRange, cosOpeningAngle=cosOpeningAngle,
                       polarizationReference=polarizationReference, stokes=stokes)
        if abs(sum(direction[k] * polarizationReference[k] for k in range(3))) > 1.0 - 1e-5:
            raise ValueError("direction and polarizationReference cannot be parallel!")
        if self._polarized and cosOpeningAngle <= 0.0:
            raise ValueError("Opening angles for polarized cone lights must be smaller 90 degrees!")

    @property
    def budget(self) -> float:
        return self._budget

    @budget.setter
    def budget(self, value: float) -> None:
        self._budget = value

    @property
    def polarized(self) -> bool:
        return self._polarized

    @property
    def _flags(self) -> int:
        return _abi.LIGHT_FLAG_POLARIZED if self._polarized else 0

    def _finishParams(self, i: int) -> None:
        super()._finishParams(i)
        c = self.budget
        self.setParam("_contribFwd", c)
        self.setParam("_contribBwd", c / (2.0 * np.pi * (1.0 - self.cosOpeningAngle)))


class PencilLightSource(LightSource):
    """Pencil beam (src/theia/light.py:1024-1102, sh/lightsource.pencil.glsl)."""

    name = "Pencil Light Source"
    _kind = _abi.LIGHT_PENCIL

    class LightParams(Structure):
        _fields_ = [("position", vec3), ("direction", vec3), ("budget", c_float), ("timeRange", vec2),
                    ("stokes", vec4), ("polarizationRef", vec3)]

    def __init__(self, *, position=(0.0, 0.0, 0.0), direction=(0.0, 0.0, 1.0), timeRange=(0.0, 100.0) * u.ns,
                 budget: float = 1.0, stokes=(1.0, 0.0, 0.0, 0.0), polarizationRef=(0.0, 1.0, 0.0)) -> None:
        super().__init__(supportForward=True, supportBackward=False, nRNGForward=1,
                         params={"LightParams": self.LightParams})
        self.setParams(position=position, direction=direction, timeRange=timeRange, budget=budget, stokes=stokes,
                       polarizationRef=polarizationRef)


class SphericalLightSource(LightSource):
    """Isotropic point source (src/theia/light.py:1105-1180, sh/lightsource.spherical.glsl)."""

    name = "Spherical Light Source"
    _kind = _abi.LIGHT_SPHERICAL

    class LightParams(Structure):
        _fields_ = [("position", vec3), ("_contribFwd", c_float), ("_contribBwd", c_float), ("timeRange", vec2)]

    def __init__(self, *, position=(0.0, 0.0, 0.0), timeRange=(0.0, 100.0) * u.ns, budget: float = 1.0) -> None:
        super().__init__(supportForward=True, supportBackward=True, nRNGForward=3, nRNGBackward=1,
                         params={"LightParams": SphericalLightSource.LightParams}, extra={"budget"})
        self.budget = budget
        self.setParams(position=position, timeRange=timeRange)

    @property
    def budget(self) -> float:
        return self._budget

    @budget.setter
    def budget(self, value: float) -> None:
        self._budget = value

    def _finishParams(self, i: int) -> None:
        super()._finishParams(i)
        c = self.budget
        self.setParam("_contribFwd", c)
        self.setParam("_contribBwd", c / (4.0 * np.pi))


class MuonTrackLightSource(LightSource):
    """
    Cherenkov light of a high energy muon track and its secondaries below 500 MeV after
    Raedel (2012) (src/theia/light.py:1413-1519, sh/lightsource.particles.muon.glsl).
    """

    name = "Muon Track Light Source"
    _kind = _abi.LIGHT_MUON

    class TrackParameters(Structure):
        _fields_ = [("startPosition", vec3), ("startTime", c_float), ("endPosition", vec3), ("endTime", c_float),
                    ("_energyScale", c_float), ("_a_angular", c_float), ("_b_angular", c_float)]

    def __init__(self, startPosition=(0.0, 0.0, 0.0), startTime: float = 0.0, endPosition=(0.0, 0.0, 0.0),
                 endTime: float = 0.0, muonEnergy: float = 1.0 * u.GeV, applyFrankTamm: bool = True) -> None:
        super().__init__(supportForward=True, nRNGForward=3, supportBackward=True, nRNGBackward=1,
                         params={"MuonTrackParams": self.TrackParameters}, extra={"muonEnergy"})
        self.setParams(startPosition=startPosition, startTime=startTime, endPosition=endPosition, endTime=endTime)
        self.muonEnergy = muonEnergy
        self._applyFrankTamm = applyFrankTamm

    @property
    def isFrankTammApplied(self) -> bool:
        return self._applyFrankTamm

    @property
    def muonEnergy(self) -> float:
        return self._muonEnergy

    @muonEnergy.setter
    def muonEnergy(self, value: float) -> None:
        self._muonEnergy = value
        # light yield scale and angular profile fits (light.py:1506-1512)
        self.setParam("_energyScale", 1.1880 + 0.0206 * np.log(value))
        self.setParam("_a_angular", 0.86634 - 7.5624e-3 * np.log10(value))
        self.setParam("_b_angular", 2.5030 + 3.0533e-2 * np.log10(value))

    @property
    def _flags(self) -> int:
        return 0 if self._applyFrankTamm else _abi.LIGHT_FLAG_FRANK_TAMM_IS


class ParticleCascadeLightSource(LightSource):
    """
    Cherenkov light of an electromagnetic or hadronic cascade after Raedel (2012):
    gamma distributed emission points along the axis, shifted exponential angular profile
    (src/theia/light.py:1522-1664, sh/lightsource.particles.cascade.glsl).
    """

    name = "Particle Cascade Light Source"
    _kind = _abi.LIGHT_CASCADE

    class CascadeParameters(Structure):
        _fields_ = [("startPosition", vec3), ("startTime", c_float), ("direction", vec3),
                    ("effectiveLength", c_float), ("a_angular", c_float), ("b_angular", c_float),
                    ("a_long", c_float), ("b_long", c_float)]

    def __init__(self, startPosition=(0.0, 0.0, 0.0), startTime: float = 0.0 * u.ns, direction=(0.0, 0.0, 1.0),
                 effectiveLength: float = 1.0, a_angular: float = 0.0, b_angular: float = 0.0, a_long: float = 0.0,
                 b_long: float = 0.0, applyFrankTamm: bool = True) -> None:
        # the gamma rejection sampler draws a data dependent number of values; like the
        # reference we reserve 12 (light.py:1629-1636)
        super().__init__(supportForward=True, supportBackward=True, nRNGForward=12, nRNGBackward=10,
                         params={"CascadeParams": self.CascadeParameters})
        self.setParams(startPosition=startPosition, startTime=startTime, direction=direction,
                       effectiveLength=effectiveLength, a_angular=a_angular, b_angular=b_angular, a_long=a_long,
                       b_long=b_long)
        self._applyFrankTamm = applyFrankTamm

    @property
    def isFrankTammApplied(self) -> bool:
        return self._applyFrankTamm

    @property
    def _flags(self) -> int:
        return 0 if self._applyFrankTamm else _abi.LIGHT_FLAG_FRANK_TAMM_IS


def frankTamm(wavelength, refractive_index):
    """
    Frank-Tamm light yield d2N/dx/dlambda in photons per (m nm) for beta = 1:
    2 pi alpha / lambda^2 * (1 - 1/n^2)  (sh/lightsource.particles.common.glsl:52-61).
    """
    lam = np.asarray(wavelength, dtype=np.float64) * 1e-3
    n = np.asarray(refractive_index, dtype=np.float64)
    return np.maximum(2.0 * np.pi * 7.2973525693 / (lam * lam) * (1.0 - 1.0 / (n * n)), 0.0)
