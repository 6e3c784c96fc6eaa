"""
Forward tracers (API of the reference's src/theia/trace.py:77-770 and 1048-1336).

A tracer gathers the committed parameter blocks of its components into one
``theia_trace_launch_t`` and hands it to ``theia_trace_forward`` (include/theia_b200.h), which
replaces the reference's shader compilation + ``dispatchIndirect``. The constructor keyword
arguments that were compile-time macros in the reference form the variant key of the launch.
"""

from __future__ import annotations

import ctypes
from abc import abstractmethod
from ctypes import Structure, addressof, byref, c_float, c_int32, c_uint32, c_uint64, memset, sizeof
from enum import IntEnum

import numpy as np
import torch
from numpy.typing import NDArray

import theia_b200.units as u
from theia_b200 import _abi, _device, _lib
from theia_b200.glsl import buffer_reference, vec3
from theia_b200.light import LightSource, WavelengthSource
from theia_b200.pipeline import Command, PipelineStage, SourceCodeMixin
from theia_b200.random import RNG, PhiloxRNG
from theia_b200.response import HitResponse, TraceConfig
from theia_b200.scene import RectBBox, Scene
from theia_b200.target import Target, TargetGuide

__all__ = [
    "EmptyEventCallback", "EventResultCode", "EventStatisticCallback", "SceneForwardTracer", "ScenePhotonTracer",
    "Tracer", "TraceEventCallback", "TrackRecordCallback", "VolumeForwardTracer", "VolumePhotonTracer",
]


class TraceEventCallback(SourceCodeMixin):
    """Base class of callbacks called for every trace event (src/theia/trace.py:49-66)."""

    name = "Trace Event Callback"
    _kind = _abi.CALLBACK_EMPTY

    def __init__(self, *, params: dict[str, type[Structure]] = {}, extra: set[str] = set()) -> None:
        super().__init__(params, extra)

    def _abiCallback(self, i: int) -> _abi.Callback:
        cb = _abi.Callback()
        cb.kind = _abi.CALLBACK_EMPTY
        return cb


class EmptyEventCallback(TraceEventCallback):
    """Ignores all events."""

    def __init__(self) -> None:
        super().__init__()


class EventStatisticCallback(TraceEventCallback):
    """
    Counts events by type (src/theia/trace.py:77-186, sh/callback.stat.glsl). Counters are
    accumulated per CTA in shared memory and flushed with one 64 bit atomic each.
    """

    name = "Event Statistic"
    _kind = _abi.CALLBACK_STAT

    class Statistic(Structure):
        _fields_ = [(n, c_uint64) for n in ("created", "scattered", "hit", "detected", "volume", "lost", "decayed",
                                            "absorbed", "missed", "maxIter", "error", "mismatch")]

    def __init__(self) -> None:
        super().__init__()
        self._tensor = torch.zeros(12, dtype=torch.int64, device=self._device)

    def _read(self) -> NDArray:
        return self._tensor.cpu().numpy()

    def __getattr__(self, name: str):
        names = [f[0] for f in EventStatisticCallback.Statistic._fields_]
        if name in names and "_tensor" in self.__dict__:
            return int(self._read()[names.index(name)])
        return super().__getattr__(name)

    def reset(self) -> None:
        self._tensor.zero_()

    def _abiCallback(self, i: int) -> _abi.Callback:
        cb = _abi.Callback()
        cb.kind = _abi.CALLBACK_STAT
        cb.stats = self._tensor.data_ptr()
        return cb

    def __str__(self) -> str:
        fields = sorted(f[0] for f in self.Statistic._fields_)
        just = max(len(f) for f in fields)
        return "\n".join(f"{f.rjust(just)}: {getattr(self, f):,}" for f in fields)


class TrackRecordCallback(TraceEventCallback):
    """
    Records the position (and optionally the polarisation state) of every event of every
    path (src/theia/trace.py:189-305, sh/callback.track.glsl).
    """

    name = "Track Recorder"
    _kind = _abi.CALLBACK_TRACK

    def __init__(self, capacity: int, length: int, *, polarized: bool = False, retrieve: bool = True) -> None:
        super().__init__()
        self._capacity = capacity
        self._length = length
        self._retrieve = retrieve
        self._polarized = polarized
        self._cols = 11 if polarized else 4
        words = capacity * length * self._cols + capacity * 2
        self._tensor = _device.alloc(words * 4, self._device)
        self._buffer = [_device.allocPinned(words * 4) for _ in range(2)]

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def length(self) -> int:
        return self._length

    @property
    def polarized(self) -> bool:
        return self._polarized

    @property
    def retrieve(self) -> bool:
        return self._retrieve

    @property
    def tensor(self) -> torch.Tensor:
        return self._tensor

    def _view(self, raw: np.ndarray) -> tuple[NDArray, NDArray, NDArray]:
        T, L, C = self.capacity, self.length, self._cols
        lengths = raw[: 4 * T].view(np.uint32)
        codes = raw[4 * T : 8 * T].view(np.int32)
        tracks = raw[8 * T : 8 * T + 4 * T * L * C].view(np.float32).reshape(C, L, T).transpose(2, 1, 0)
        return tracks, lengths, codes

    def result(self, i: int) -> tuple[NDArray, NDArray, NDArray]:
        """(tracks[ray, event, coord], lengths[ray], codes[ray]) of configuration i"""
        return self._view(self._buffer[i].numpy())

    def _abiCallback(self, i: int) -> _abi.Callback:
        cb = _abi.Callback()
        cb.kind = _abi.CALLBACK_TRACK
        cb.track_polarized = int(self.polarized)
        cb.track_buffer = self._tensor.data_ptr()
        cb.track_count = self.capacity
        cb.track_length = self.length
        return cb

    def run(self, i: int) -> list[Command]:
        if not self.retrieve:
            return []
        return [lambda: self._buffer[i].copy_(self._tensor, non_blocking=True)]


class EventResultCode(IntEnum):
    """Result codes passed to callbacks; negative codes stop the path (trace.py:308-343)."""

    SUCCESS = 0
    RAY_CREATED = 1
    RAY_SCATTERED = 2
    RAY_HIT = 3
    RAY_DETECTED = 4
    VOLUME_HIT = 5
    RAY_LOST = -1
    RAY_DECAYED = -2
    RAY_ABSORBED = -3
    RAY_MISSED = -4
    MAX_ITER = -5
    ERROR_CODE_MAX_VALUE = -10
    ERROR_UNKNOWN = -10
    ERROR_MEDIA_MISMATCH = -11
    ERROR_TRACE_ABORT = -12
    ERROR_RAY_BAD = -13


def _component(stage: SourceCodeMixin | None, i: int, keep: list) -> _abi.Component:
    c = _abi.Component()
    if stage is None:
        return c
    block = stage._firstBlock(i)
    c.kind = stage._kind
    c.flags = stage._flags
    if block is not None:
        keep.append(block)
        c.params = addressof(block)
        c.params_size = sizeof(block)
    return c


class Tracer(PipelineStage):
    """Base class of tracers (src/theia/trace.py:346-497)."""

    name = "Tracer"
    _tracerKind = 0

    class DispatchParams(Structure):
        _fields_ = [("_batchSize", c_uint32), ("_groupCountX", c_uint32), ("_groupCountY", c_uint32),
                    ("_groupCountZ", c_uint32)]

    def __init__(self, response: HitResponse, params: dict[str, type[Structure]] = {}, extra: set[str] = set(), *,
                 batchSize: int, blockSize: int, capacity: int | None, nRNGSamples: int, maxHitsPerThread: int,
                 polarized: bool = False) -> None:
        params = {**params, "DispatchParams": self.DispatchParams}
        extra = {*extra, "batchSize"}
        super().__init__(params, extra)
        self._blockSize = blockSize
        self._capacity = batchSize if capacity is None else capacity
        self._response = response
        self._maxHitsPerThread = maxHitsPerThread
        self._nRNGSamples = nRNGSamples
        self._polarized = polarized
        self._firstPath = 0
        self._workCounters = None
        self._rowCache: dict = {}
        self.setParams(batchSize=batchSize, _groupCountY=1, _groupCountZ=1)
        response.prepare(self._getTraceConfig())

    @property
    def batchSize(self) -> int:
        """Number of paths per run"""
        return self.getParam("_batchSize")

    @batchSize.setter
    def batchSize(self, value: int) -> None:
        if value < 0:
            raise ValueError("batchSize must be at least zero!")
        if value > self.capacity:
            raise ValueError("batchSize cannot be larger than capacity!")
        groupCount = max(-(value // -self.blockSize), 1)
        self.setParams(_batchSize=value, _groupCountX=groupCount)

    @property
    def blockSize(self) -> int:
        return self._blockSize

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def maxHits(self) -> int:
        return self.maxHitsPerThread * self.batchSize

    @property
    def maxHitsPerThread(self) -> int:
        return self._maxHitsPerThread

    @property
    def normalization(self) -> float:
        return 1.0 / self.batchSize

    @property
    def nRNGSamples(self) -> int:
        return self._nRNGSamples

    @property
    def polarized(self) -> bool:
        return self._polarized

    @property
    def response(self) -> HitResponse:
        return self._response

    @property
    def firstPath(self) -> int:
        """
        B200 extension: global index of the first path (RNG stream) of this tracer's batches;
        lets several GPUs split one batch. Default 0.
        """
        return self._firstPath

    @firstPath.setter
    def firstPath(self, value: int) -> None:
        self._firstPath = int(value)

    def enableWorkCounters(self) -> torch.Tensor:
        """
        B200 extension (measurement): device counters every launch adds to, int64[4] =
        path segments (the sum of the reference's EventStatisticCallback counters), hits,
        shadow rays, paths started. Returns the tensor.
        """
        if self._workCounters is None:
            self._workCounters = torch.zeros(4, dtype=torch.int64, device=self._device)
        return self._workCounters

    def collectStages(self, *, useOriginalNames: bool = False) -> list[tuple[str, PipelineStage]]:
        stages = self._collectStages()
        if useOriginalNames:
            return [(s.name, s) for _, s in stages]
        return stages

    @abstractmethod
    def _collectStages(self) -> list[tuple[str, PipelineStage]]:
        pass

    def _getTraceConfig(self) -> TraceConfig:
        return TraceConfig(self.batchSize, self.blockSize, self.capacity, self.maxHitsPerThread, self.normalization,
                           self.polarized)

    def _finishParams(self, i: int) -> None:
        super()._finishParams(i)
        self.response.updateConfig(self._getTraceConfig())

    # ---- launch assembly
    def _fillLaunch(self, L: _abi.TraceLaunch, i: int, keep: list) -> None:
        raise NotImplementedError

    def _launch(self, i: int) -> tuple[_abi.TraceLaunch, list]:
        """Assembles the C-ABI launch description from the parameters committed to config i."""
        keep: list = []
        L = _abi.TraceLaunch()
        L.struct_size = sizeof(_abi.TraceLaunch)
        L.tracer = self._tracerKind
        L.polarized = int(self.polarized)
        L.block_size = self.blockSize
        L.batch_size = self._block("DispatchParams", i)._batchSize
        L.first_path = self._firstPath
        trace = self._block("TraceParams", i)
        keep.append(trace)
        L.trace_params = addressof(trace)
        L.trace_params_size = sizeof(trace)
        if not isinstance(self.rng, PhiloxRNG):
            raise RuntimeError("only PhiloxRNG is available in this build")
        L.rng = self.rng._abiParams(i)
        L.light = _component(self.source, i, keep)
        L.wavelength = _component(self.wavelengthSource, i, keep)
        self._fillLaunch(L, i, keep)
        nBlocks = -(L.batch_size // -self.blockSize)
        L.response = self.response._abiResponse(i, nBlocks)
        L.callback = self.callback._abiCallback(i)
        L.stream = _device.currentStream(self._device)
        if self._workCounters is not None:
            L.work_counters = self._workCounters.data_ptr()
        if L.response.kind == _abi.RESPONSE_HISTOGRAM:
            # only the rows the launch really writes are summed (and cleared) by the reducer
            rows = min(self._histRows(L), L.response.hist_capacity)
            L.response.hist_capacity = rows
            self.response._setRows(i, rows)
        return L, keep

    def _histRows(self, L: _abi.TraceLaunch) -> int:
        """Number of partial histograms a launch writes: one per CTA of its (persistent) grid."""
        if self._device != "cuda":
            return -(L.batch_size // -self.blockSize)
        key = (L.batch_size, bool(L.workspace), L.polarized, L.response.n_bins)
        rows = self._rowCache.get(key)
        if rows is None:
            g = c_uint32()
            _lib.check(_lib.load().theia_trace_grid_size(byref(L), byref(g)))
            rows = self._rowCache[key] = int(g.value)
        return rows

    def run(self, i: int) -> list[Command]:
        def cmd() -> None:
            if self._device != "cuda":
                raise RuntimeError(f"{self.name} needs a CUDA device: theia_b200 has no CPU fallback")
            L, keep = self._launch(i)
            _lib.check(_lib.load().theia_trace_forward(byref(L)))

        return [cmd]


class VolumeForwardTracer(Tracer):
    """
    Forward path tracer through a homogeneous medium towards a ``Target``, without scene
    intersections (src/theia/trace.py:499-770, sh/tracer.volume.forward.glsl).

    Parameters follow the reference: ``batchSize, source, target, wavelengthSource, response,
    rng, *, medium, objectId=0, capacity=None, callback, nScattering=6, scatterCoefficient=NaN,
    traceBBox, maxTime=1000 ns, polarized=False, disableDirectLighting=False,
    disableTargetSampling=False, blockSize=128``. Stage parameters: ``scatterCoefficient,
    medium, objectId, lowerBBoxCorner, upperBBoxCorner, maxTime``.
    """

    name = "Volume Forward Tracer"
    _tracerKind = _abi.TRACER_VOLUME_FORWARD

    class TraceParams(Structure):
        _pack_ = 4
        _fields_ = [("medium", buffer_reference), ("objectId", c_int32), ("scatterCoefficient", c_float),
                    ("lowerBBoxCorner", vec3), ("upperBBoxCorner", vec3), ("maxTime", c_float), ("_maxDist", c_float)]

    def __init__(self, batchSize: int, source: LightSource, target: Target, wavelengthSource: WavelengthSource,
                 response: HitResponse, rng: RNG, *, medium: int, objectId: int = 0, capacity: int | None = None,
                 callback: TraceEventCallback | None = None, nScattering: int = 6,
                 scatterCoefficient: float = float("NaN"), traceBBox: RectBBox | None = None,
                 maxTime: float = 1000.0 * u.ns, polarized: bool = False, disableDirectLighting: bool = False,
                 disableTargetSampling: bool = False, blockSize: int = 128, code: bytes | None = None) -> None:
        if callback is None:
            callback = EmptyEventCallback()
        if traceBBox is None:
            traceBBox = RectBBox((-1.0 * u.km,) * 3, (1.0 * u.km,) * 3)
        # budgets (trace.py:623-634)
        maxHitsPerThread = nScattering
        if not disableTargetSampling:
            maxHitsPerThread *= 2
        if not disableDirectLighting:
            maxHitsPerThread += 1
        pathLength = nScattering if disableTargetSampling else nScattering - 1
        rngStride = 3 if disableTargetSampling else 7
        nRNG = source.nRNGForward + wavelengthSource.nRNGSamples
        nRNG += rngStride * pathLength
        nRNG += maxHitsPerThread * response.nRNGSamples
        super().__init__(response, {"TraceParams": self.TraceParams}, batchSize=batchSize, blockSize=blockSize,
                         capacity=capacity, nRNGSamples=nRNG, maxHitsPerThread=maxHitsPerThread, polarized=polarized)
        self._source = source
        self._wavelengthSource = wavelengthSource
        self._rng = rng
        self._target = target
        self._callback = callback
        self._nScattering = nScattering
        self._pathLength = pathLength
        self._directLightingDisabled = disableDirectLighting
        self._targetSamplingDisabled = disableTargetSampling
        self.setParams(scatterCoefficient=scatterCoefficient, medium=medium, objectId=objectId,
                       lowerBBoxCorner=traceBBox.lowerCorner, upperBBoxCorner=traceBBox.upperCorner, maxTime=maxTime,
                       _maxDist=traceBBox.diagonal)

    @property
    def callback(self) -> TraceEventCallback:
        return self._callback

    @property
    def code(self) -> bytes:
        """Variant key (the reference returns SPIR-V here)"""
        return repr((self.name, self.polarized, self._directLightingDisabled, self._targetSamplingDisabled,
                     self._pathLength, self.blockSize)).encode()

    @property
    def directLightingDisabled(self) -> bool:
        return self._directLightingDisabled

    @property
    def nScattering(self) -> int:
        return self._nScattering

    @property
    def rng(self) -> RNG:
        return self._rng

    @property
    def source(self) -> LightSource:
        return self._source

    @property
    def target(self) -> Target:
        return self._target

    @property
    def targetSamplingDisabled(self) -> bool:
        return self._targetSamplingDisabled

    @property
    def traceBBox(self) -> RectBBox:
        return RectBBox(self.getParam("lowerBBoxCorner"), self.getParam("upperBBoxCorner"))

    @traceBBox.setter
    def traceBBox(self, value: RectBBox) -> None:
        self.setParams(lowerBBoxCorner=value.lowerCorner, upperBBoxCorner=value.upperCorner, _maxDist=value.diagonal)

    @property
    def wavelengthSource(self) -> WavelengthSource:
        return self._wavelengthSource

    def _collectStages(self) -> list[tuple[str, PipelineStage]]:
        return [("rng", self.rng), ("photons", self.wavelengthSource), ("lightSource", self.source),
                ("target", self.target), ("tracer", self), ("callback", self.callback), ("response", self.response)]

    def _finishParams(self, i: int) -> None:
        super()._finishParams(i)
        self.setParam("_maxDist", self.traceBBox.diagonal)

    def _fillLaunch(self, L: _abi.TraceLaunch, i: int, keep: list) -> None:
        L.disable_mis = int(self._targetSamplingDisabled)
        L.disable_direct_lighting = int(self._directLightingDisabled)
        L.path_length = self._pathLength
        L.target = _component(self.target, i, keep)


class SceneForwardTracer(Tracer):
    """
    Forward path tracer through a meshed scene with volume scattering, Fresnel reflection /
    transmission, volume borders, detectors and optional next event estimation towards a
    ``TargetGuide`` (src/theia/trace.py:1048-1336, sh/tracer.scene.forward.glsl,
    sh/scene.traverse.glsl).

    Parameters follow the reference: ``batchSize, source, wavelengthSource, response, rng,
    scene, *, capacity=None, callback, maxPathLength=6, targetId=-1, targetGuide=None,
    scatterCoefficient=NaN, sourceMedium=None, maxTime=1000 ns, polarized=False,
    disableDirectLighting=False, disableTransmission=False, disableVolumeBorder=False,
    useRefractedHitDir=False, blockSize=128``. Stage parameters: ``targetId,
    scatterCoefficient, maxTime``.
    """

    name = "Scene Forward Tracer"
    _tracerKind = _abi.TRACER_SCENE_FORWARD

    class TraceParams(Structure):
        _pack_ = 4
        _fields_ = [("targetId", c_int32), ("sourceMedium", buffer_reference), ("scatterCoefficient", c_float),
                    ("_lowerBBoxCorner", vec3), ("_upperBBoxCorner", vec3), ("maxTime", c_float),
                    ("_maxDist", c_float)]

    def __init__(self, batchSize: int, source: LightSource, wavelengthSource: WavelengthSource,
                 response: HitResponse, rng: RNG, scene: Scene, *, capacity: int | None = None,
                 callback: TraceEventCallback | None = None, maxPathLength: int = 6, targetId: int = -1,
                 targetGuide: TargetGuide | None = None, scatterCoefficient: float = float("NaN"),
                 sourceMedium: int | None = None, maxTime: float = 1000.0 * u.ns, polarized: bool = False,
                 disableDirectLighting: bool = False, disableTransmission: bool = False,
                 disableVolumeBorder: bool = False, useRefractedHitDir: bool = False, blockSize: int = 128,
                 code: bytes | None = None, mode: str = "auto") -> None:
        if callback is None:
            callback = EmptyEventCallback()
        if mode not in ("auto", "wavefront", "single"):
            raise ValueError("mode must be 'auto', 'wavefront' or 'single'")
        # B200 extension: "wavefront" keeps the path state in HBM and runs one trace + one
        # shade kernel per path segment; "single" runs the whole batch in one launch with the
        # paths resident in registers (small batches). Results are identical.
        self._mode = mode
        self._workspace = None
        # budgets (trace.py:1176-1187)
        maxHits = maxPathLength - 1
        rngStride = 4
        if targetGuide is not None:
            maxHits *= 2
            rngStride += targetGuide.nRNGSamples
        if not disableDirectLighting:
            maxHits += 1
        nRNG = source.nRNGForward + wavelengthSource.nRNGSamples
        nRNG += rngStride * maxPathLength
        nRNG += maxHits * response.nRNGSamples
        super().__init__(response, {"TraceParams": self.TraceParams}, batchSize=batchSize, blockSize=blockSize,
                         capacity=capacity, maxHitsPerThread=maxHits, nRNGSamples=nRNG, polarized=polarized)
        if sourceMedium is None:
            sourceMedium = scene.medium
        self._source = source
        self._wavelengthSource = wavelengthSource
        self._callback = callback
        self._rng = rng
        self._scene = scene
        self._targetGuide = targetGuide
        self._maxPathLength = maxPathLength
        self._directLightingDisabled = disableDirectLighting
        self._transmissionDisabled = disableTransmission
        self._volumeBorderDisabled = disableVolumeBorder
        self._useRefractedHitDir = useRefractedHitDir
        self.setParams(targetId=targetId, scatterCoefficient=scatterCoefficient, sourceMedium=sourceMedium,
                       maxTime=maxTime, _lowerBBoxCorner=scene.bbox.lowerCorner,
                       _upperBBoxCorner=scene.bbox.upperCorner, _maxDist=scene.bbox.diagonal)

    @property
    def callback(self) -> TraceEventCallback:
        return self._callback

    @property
    def code(self) -> bytes:
        return repr((self.name, self.polarized, self._directLightingDisabled, self._targetGuide is None,
                     self._transmissionDisabled, self._volumeBorderDisabled, self._useRefractedHitDir,
                     self._maxPathLength, self.blockSize)).encode()

    @property
    def directLightingDisabled(self) -> bool:
        return self._directLightingDisabled

    @property
    def hitsUseRefractedDir(self) -> bool:
        return self._useRefractedHitDir

    @property
    def maxPathLength(self) -> int:
        return self._maxPathLength

    @property
    def rng(self) -> RNG:
        return self._rng

    @property
    def scene(self) -> Scene:
        return self._scene

    @property
    def source(self) -> LightSource:
        return self._source

    @property
    def targetGuide(self) -> TargetGuide | None:
        return self._targetGuide

    @property
    def transmissionDisabled(self) -> bool:
        return self._transmissionDisabled

    @property
    def volumeBorderDisabled(self) -> bool:
        return self._volumeBorderDisabled

    @property
    def wavelengthSource(self) -> WavelengthSource:
        return self._wavelengthSource

    def _collectStages(self) -> list[tuple[str, PipelineStage]]:
        stages = [("rng", self.rng), ("photons", self.wavelengthSource), ("lightSource", self.source)]
        if self.targetGuide is not None:
            stages.append(("guide", self.targetGuide))
        stages += [("tracer", self), ("callback", self.callback), ("response", self.response)]
        return stages

    def _fillLaunch(self, L: _abi.TraceLaunch, i: int, keep: list) -> None:
        L.disable_mis = int(self._targetGuide is None)
        L.disable_direct_lighting = int(self._directLightingDisabled)
        L.disable_transmission = int(self._transmissionDisabled)
        L.disable_volume_border = int(self._volumeBorderDisabled)
        L.hit_use_transmitted_dir = int(self._useRefractedHitDir)
        L.path_length = self._maxPathLength
        L.guide = _component(self._targetGuide, i, keep)
        tlas = self._scene.tlas
        L.tlas = 0 if tlas is None else tlas.data_ptr()
        if self.wavefront and self._device == "cuda":
            need = ctypes.c_uint64()
            _lib.check(_lib.load().theia_trace_workspace_size(byref(L), byref(need)))
            if self._workspace is None or self._workspace.numel() < need.value:
                self._workspace = _device.alloc(need.value, "cuda", zero=False)
            L.workspace = self._workspace.data_ptr()
            L.workspace_bytes = self._workspace.numel()

    @property
    def wavefront(self) -> bool:
        """True if batches run in wavefront form (B200 extension, see ``mode``)"""
        if self._mode == "auto":
            return self.capacity >= 65536
        return self._mode == "wavefront"


class VolumePhotonTracer(VolumeForwardTracer):
    """
    Traces single photons through a homogeneous medium towards a ``Target``: every scattering
    event plays roulette on the survival probability instead of carrying a weight, and a photon
    that reaches the target is handed to the response once and absorbed
    (src/theia/trace.py:2657-2959, sh/tracer.volume.photon.*.glsl). Pair it with
    ``StoreTimeHitResponse`` to obtain detected photons.

    Parameters follow the reference: ``batchSize, source, target, wavelengthSource, response, rng,
    *, medium, objectId=0, capacity=None, callback, nScatteringPerRun=10, nRuns=10, traceBBox,
    maxTime=1 ms, polarized=False, blockSize=128``. The reference dispatches one shader per run
    and hands the surviving photons over in a queue; here a photon simply lives on in the
    registers of its lane, with the draw counter of run r starting at ``r * nRNGPerRun`` as in
    the reference.
    """

    name = "Volume Photon Tracer"
    _tracerKind = _abi.TRACER_VOLUME_PHOTON

    def __init__(self, batchSize: int, source: LightSource, target: Target, wavelengthSource: WavelengthSource,
                 response: HitResponse, rng: RNG, *, medium: int, objectId: int = 0, capacity: int | None = None,
                 callback: TraceEventCallback | None = None, nScatteringPerRun: int = 10, nRuns: int = 10,
                 traceBBox: RectBBox | None = None, maxTime: float = 1.0 * u.ms, polarized: bool = False,
                 blockSize: int = 128, code: bytes | None = None) -> None:
        if callback is None:
            callback = EmptyEventCallback()
        if traceBBox is None:
            traceBBox = RectBBox((-1.0 * u.km,) * 3, (1.0 * u.km,) * 3)
        if nScatteringPerRun <= 0 or nRuns <= 0:
            raise ValueError("nScatteringPerRun and nRuns must be positive")
        # budgets (trace.py:2784-2787)
        nRNG = source.nRNGForward + wavelengthSource.nRNGSamples
        nRNG += 4 * nRuns * nScatteringPerRun
        nRNG += response.nRNGSamples
        self._nRNGPerRun = 4 * nScatteringPerRun + response.nRNGSamples
        Tracer.__init__(self, response, {"TraceParams": self.TraceParams}, batchSize=batchSize, blockSize=blockSize,
                        capacity=capacity, nRNGSamples=nRNG, maxHitsPerThread=1, polarized=polarized)
        self._source = source
        self._wavelengthSource = wavelengthSource
        self._rng = rng
        self._target = target
        self._callback = callback
        self._nRuns = nRuns
        self._nScatteringPerRun = nScatteringPerRun
        self._nScattering = nRuns * nScatteringPerRun
        self._pathLength = nRuns * nScatteringPerRun
        self._directLightingDisabled = False
        self._targetSamplingDisabled = True
        self.setParams(scatterCoefficient=float("NaN"), medium=medium, objectId=objectId,
                       lowerBBoxCorner=traceBBox.lowerCorner, upperBBoxCorner=traceBBox.upperCorner, maxTime=maxTime,
                       _maxDist=traceBBox.diagonal)

    @property
    def nScatteringPerRun(self) -> int:
        """Scattering events per run of the reference's loop shader"""
        return self._nScatteringPerRun

    @property
    def nRuns(self) -> int:
        return self._nRuns

    @property
    def maxPathLength(self) -> int:
        return self._nRuns * self._nScatteringPerRun

    @property
    def code(self) -> bytes:
        return repr((self.name, self.polarized, self._nRuns, self._nScatteringPerRun, self.blockSize)).encode()

    def _fillLaunch(self, L: _abi.TraceLaunch, i: int, keep: list) -> None:
        super()._fillLaunch(L, i, keep)
        L.path_length = self.maxPathLength
        L.run_length = self._nScatteringPerRun
        L.rng_per_run = self._nRNGPerRun


class ScenePhotonTracer(SceneForwardTracer):
    """
    Traces single photons through a meshed scene with absorption roulette after every
    interaction (src/theia/trace.py:2370-2654, sh/tracer.scene.photon.*.glsl); no next event
    estimation, every detector hit goes to the response. Parameters follow the reference:
    ``batchSize, source, wavelengthSource, response, rng, scene, *, capacity=None, callback,
    targetId=-1, sourceMedium=None, maxTime=1 ms, nScatteringPerRun=10, nRuns=10,
    polarized=False, disableTransmission=False, disableVolumeBorder=False,
    useRefractedHitDir=False, blockSize=128`` (+ ``mode`` as for ``SceneForwardTracer``).
    """

    name = "Scene Photon Tracer"
    _tracerKind = _abi.TRACER_SCENE_PHOTON

    def __init__(self, batchSize: int, source: LightSource, wavelengthSource: WavelengthSource,
                 response: HitResponse, rng: RNG, scene: Scene, *, capacity: int | None = None,
                 callback: TraceEventCallback | None = None, targetId: int = -1, sourceMedium: int | None = None,
                 maxTime: float = 1.0 * u.ms, nScatteringPerRun: int = 10, nRuns: int = 10, polarized: bool = False,
                 disableTransmission: bool = False, disableVolumeBorder: bool = False,
                 useRefractedHitDir: bool = False, blockSize: int = 128, code: bytes | None = None,
                 mode: str = "auto") -> None:
        if callback is None:
            callback = EmptyEventCallback()
        if mode not in ("auto", "wavefront", "single"):
            raise ValueError("mode must be 'auto', 'wavefront' or 'single'")
        if nScatteringPerRun <= 0 or nRuns <= 0:
            raise ValueError("nScatteringPerRun and nRuns must be positive")
        self._mode = mode
        self._workspace = None
        # budgets (trace.py:2488-2491)
        nRNG = source.nRNGForward + wavelengthSource.nRNGSamples
        nRNG += 5 * nRuns * nScatteringPerRun
        nRNG += response.nRNGSamples
        self._nRNGPerRun = 5 * nScatteringPerRun + response.nRNGSamples
        Tracer.__init__(self, response, {"TraceParams": self.TraceParams}, batchSize=batchSize, blockSize=blockSize,
                        capacity=capacity, maxHitsPerThread=1, nRNGSamples=nRNG, polarized=polarized)
        if sourceMedium is None:
            sourceMedium = scene.medium
        self._source = source
        self._wavelengthSource = wavelengthSource
        self._callback = callback
        self._rng = rng
        self._scene = scene
        self._targetGuide = None
        self._nRuns = nRuns
        self._nScatteringPerRun = nScatteringPerRun
        self._maxPathLength = nRuns * nScatteringPerRun
        self._directLightingDisabled = False
        self._transmissionDisabled = disableTransmission
        self._volumeBorderDisabled = disableVolumeBorder
        self._useRefractedHitDir = useRefractedHitDir
        self.setParams(targetId=targetId, scatterCoefficient=float("NaN"), sourceMedium=sourceMedium,
                       maxTime=maxTime, _lowerBBoxCorner=scene.bbox.lowerCorner,
                       _upperBBoxCorner=scene.bbox.upperCorner, _maxDist=scene.bbox.diagonal)

    @property
    def nScatteringPerRun(self) -> int:
        return self._nScatteringPerRun

    @property
    def nRuns(self) -> int:
        return self._nRuns

    @property
    def code(self) -> bytes:
        return repr((self.name, self.polarized, self._transmissionDisabled, self._volumeBorderDisabled,
                     self._useRefractedHitDir, self._nRuns, self._nScatteringPerRun, self.blockSize)).encode()

    def _fillLaunch(self, L: _abi.TraceLaunch, i: int, keep: list) -> None:
        super()._fillLaunch(L, i, keep)
        L.run_length = self._nScatteringPerRun
        L.rng_per_run = self._nRNGPerRun
