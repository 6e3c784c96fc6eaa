"""
Hit responses (API of the reference's src/theia/response.py, forward path): what happens
with a hit once a tracer has produced it - recorded into a queue or binned into a time
histogram ("light curve").
"""

from __future__ import annotations

import ctypes
from ctypes import Structure, byref, c_float, c_int32, c_uint32
from dataclasses import dataclass
from typing import Final
from warnings import warn

import numpy as np
import torch
from numpy.typing import NDArray

import theia_b200.units as u
from theia_b200 import _abi, _device, _lib
from theia_b200.pipeline import Command, PipelineStage, SourceCodeMixin
from theia_b200.queue import IOQueue
from theia_b200.random import RNG

__all__ = [
    "EmptyResponse", "HistogramHitResponse", "HistogramReducer", "HitItem", "HitRecorder", "HitReplay",
    "HitResponse", "HitTimeAndIdItem", "HitTimeItem", "KernelHistogramHitResponse", "ObjectHistogramHitResponse",
    "PolarizedHitItem", "StoreTimeHitResponse", "TraceConfig", "UniformValueResponse", "ValueItem", "ValueResponse",
]


class HitItem(Structure):
    """Layout of one hit (src/theia/response.py:55-70, sh/response.common.glsl:4-19)."""

    _fields_ = [
        ("position", c_float * 3), ("direction", c_float * 3), ("normal", c_float * 3),
        ("wavelength", c_float), ("time", c_float), ("contrib", c_float), ("objectId", c_int32),
    ]


class PolarizedHitItem(Structure):
    """Hit with Stokes vector and polarisation reference (src/theia/response.py:73-92)."""

    _fields_ = [
        ("position", c_float * 3), ("direction", c_float * 3), ("normal", c_float * 3),
        ("stokes", c_float * 4), ("polarizationRef", c_float * 3),
        ("wavelength", c_float), ("time", c_float), ("contrib", c_float), ("objectId", c_int32),
    ]


class ValueItem(Structure):
    _fields_ = [("value", c_float), ("time", c_float)]


class HitTimeItem(Structure):
    """Item of ``StoreTimeHitResponse`` storing only the time (src/theia/response.py:626-629)"""

    _fields_ = [("time", c_float)]


class HitTimeAndIdItem(Structure):
    """Item of ``StoreTimeHitResponse`` storing time and object id (src/theia/response.py:632-635)"""

    _fields_ = [("time", c_float), ("objectId", c_int32)]


@dataclass
class TraceConfig:
    """What a tracer tells its response about itself (src/theia/response.py:95-122)."""

    batchSize: int
    blockSize: int
    capacity: int
    maxHitsPerThread: int = 1
    normalization: float = 1.0
    polarized: bool = False


class HitResponse(SourceCodeMixin):
    """
    Base class of responses, called by a tracer for every hit (src/theia/response.py:125-178).
    ``prepare`` is called once per tracer at construction, ``updateConfig`` before every batch.
    """

    name = "Hit Response"
    _kind = _abi.RESPONSE_EMPTY

    def __init__(self, params: dict[str, type[Structure]] = {}, extra: set[str] = set(), *,
                 nRNGSamples: int = 0) -> None:
        super().__init__(params, extra)
        self._nRNGSamples = nRNGSamples

    @property
    def nRNGSamples(self) -> int:
        return self._nRNGSamples

    def prepare(self, config: TraceConfig) -> None:
        pass

    def updateConfig(self, config: TraceConfig) -> None:
        pass

    def _abiResponse(self, i: int, nBlocks: int) -> _abi.Response:
        """C-ABI description of this response for configuration i"""
        r = _abi.Response()
        r.kind = _abi.RESPONSE_EMPTY
        return r

    def _partialCapacity(self) -> int:
        return 0


class EmptyResponse(HitResponse):
    """Ignores all hits."""

    def __init__(self) -> None:
        super().__init__()


class HitRecorder(HitResponse):
    """
    Records hits into a queue that is copied to the host after each batch
    (src/theia/response.py:191-275, sh/response.record.glsl). Slots are handed out by a
    warp-aggregated atomic counter, so the order of hits is not defined.
    """

    name = "Hit Recorder"
    _kind = _abi.RESPONSE_RECORD

    def __init__(self, *, capacity: int = 0, polarized: bool = False) -> None:
        super().__init__()
        self._capacity = capacity
        self._polarized = polarized
        self._queue = IOQueue(PolarizedHitItem if polarized else HitItem, mode="retrieve")
        if capacity > 0:
            self._queue.initialize(capacity)
        self._tags = None

    def prepare(self, config: TraceConfig) -> None:
        maxHits = config.maxHitsPerThread * config.capacity
        if self.queue.initialized and maxHits > self.capacity:
            warn(f"tracer reported maxHits: {maxHits}, but response was initialized with capacity: {self.capacity}")
        elif not self.queue.initialized:
            self._capacity = max(self._capacity, maxHits)

    def updateConfig(self, config: TraceConfig) -> None:
        maxHits = config.batchSize * config.maxHitsPerThread
        if not self.queue.initialized:
            self._capacity = max(maxHits, self._capacity)
        elif maxHits > self.capacity:
            warn(f"tracer reported maxHits: {maxHits}, but response was initialized with capacity: {self.capacity}")

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def polarized(self) -> bool:
        return self._polarized

    @property
    def queue(self) -> IOQueue:
        return self._queue

    def enableTags(self) -> None:
        """
        B200 extension (test seam): additionally records which path and which event
        produced each hit; see ``tags``.
        """
        self._wantTags = True

    def tags(self) -> tuple[NDArray, NDArray]:
        """(path index, event number) of every queue slot, valid up to the hit count."""
        return self._tags[0].cpu().numpy(), self._tags[1].cpu().numpy()

    def _ensure(self) -> None:
        if not self.queue.initialized:
            if self.capacity <= 0:
                raise RuntimeError(f"Cannot create a queue of size {self.capacity}")
            self.queue.initialize(self.capacity)
        if getattr(self, "_wantTags", False) and self._tags is None:
            self._tags = [torch.zeros(self.capacity, dtype=torch.int32, device=self._device) for _ in range(2)]

    def _abiResponse(self, i: int, nBlocks: int) -> _abi.Response:
        self._ensure()
        r = _abi.Response()
        r.kind = _abi.RESPONSE_RECORD
        r.queue = self.queue.tensor.data_ptr()
        r.queue_capacity = self.capacity
        r.queue_polarized = int(self.polarized)
        if self._tags is not None:
            r.tag_path = self._tags[0].data_ptr()
            r.tag_event = self._tags[1].data_ptr()
        return r

    def run(self, i: int) -> list[Command]:
        self._ensure()
        return self.queue.run(i)


class ValueResponse(SourceCodeMixin):
    """Base class of functions turning a hit into a number (src/theia/response.py:444-483)."""

    name = "Value Response"
    _kind = 0

    def __init__(self, params: dict[str, type[Structure]] = {}, extra: set[str] = set(), *,
                 nRNGSamples: int = 0) -> None:
        super().__init__(params, extra)
        self._nRNGSamples = nRNGSamples

    @property
    def nRNGSamples(self) -> int:
        return self._nRNGSamples

    def prepare(self, config: TraceConfig) -> None:
        pass


class UniformValueResponse(ValueResponse):
    """Perfect isotropic detector: value = hit contribution (sh/response.uniform.glsl:4-6)."""

    _kind = _abi.VALUE_UNIFORM

    def __init__(self) -> None:
        super().__init__()


class HistogramReducer(PipelineStage):
    """
    Reduces partial histograms into one: per bin a Kahan sum over the partials times
    ``normalization`` (src/theia/response.py:1065-1197, sh/estimator.reduce.glsl). The
    partials are cleared afterwards.
    """

    name = "Histogram Reducer"
    MAX_SHARED_MEMORY_BINS: Final[int] = 8192

    class Params(Structure):
        _fields_ = [("normalization", c_float), ("nHist", c_uint32)]

    def __init__(self, *, nBins: int, nHist: int, histIn: torch.Tensor | None = None,
                 histOut: torch.Tensor | None = None, retrieve: bool = True, normalization: float = 1.0,
                 blockSize: int = 128, useSharedMemory: bool = True) -> None:
        super().__init__({"Params": self.Params})
        self._blockSize = blockSize
        self._nBins = nBins
        self._retrieve = retrieve
        self.setParams(normalization=normalization, nHist=nHist)
        self._useSharedMemory = useSharedMemory
        dev = self._device
        self._histIn = torch.zeros(nBins * nHist, dtype=torch.float32, device=dev) if histIn is None else histIn
        self._histOut = torch.zeros(nBins, dtype=torch.float32, device=dev) if histOut is None else histOut
        pin = torch.cuda.is_available()
        self._buffer = [torch.zeros(nBins, dtype=torch.float32, pin_memory=pin) if retrieve else None for _ in range(2)]

    @property
    def blockSize(self) -> int:
        return self._blockSize

    @property
    def capacity(self) -> int:
        """Maximum number of partial histograms"""
        return self._histIn.numel() // self.nBins

    @property
    def histIn(self) -> torch.Tensor:
        return self._histIn

    @property
    def histOut(self) -> torch.Tensor:
        return self._histOut

    @property
    def nBins(self) -> int:
        return self._nBins

    @property
    def retrieve(self) -> bool:
        return self._retrieve

    def result(self, i: int) -> NDArray | None:
        return self._buffer[i].numpy() if self.retrieve else None

    def run(self, i: int) -> list[Command]:
        def cmd() -> None:
            if self._device != "cuda":
                raise RuntimeError("HistogramReducer needs a CUDA device (no CPU fallback)")
            p = self._block("Params", i)
            _lib.check(_lib.load().theia_hist_reduce(
                self._histIn.data_ptr(), p.nHist, self.nBins, p.normalization, self._histOut.data_ptr(), 1,
                _device.currentStream("cuda")))
            if self.retrieve:
                self._buffer[i].copy_(self._histOut, non_blocking=True)

        return [cmd]


class HistogramHitResponse(HitResponse):
    """
    Bins the values of hits by arrival time (src/theia/response.py:1200-1421,
    sh/response.histogram.glsl). Every CTA of the tracer accumulates a partial histogram in
    shared memory with warp-aggregated float atomics and writes it out once; a reduce stage
    sums the partials.

    Parameters: ``response`` (value response), ``nBins``, ``t0`` (first bin edge), ``binSize``,
    ``normalization`` (None: use the tracer's 1 / batchSize), ``retrieve``.
    """

    name = "Histogram Hit Response"
    _kind = _abi.RESPONSE_HISTOGRAM

    class Params(Structure):
        _fields_ = [("t0", c_float), ("binSize", c_float)]

    def __init__(self, response: ValueResponse, *, nBins: int = 100, t0: float = 0.0 * u.ns,
                 binSize: float = 1.0 * u.ns, normalization: float | None = None, retrieve: bool = True,
                 reduceBlockSize: int = 128, updateResponse: bool = True, useSharedMemory: bool = True) -> None:
        HitResponse.__init__(self, {"ResponseParams": self.Params}, {"normalization"}, nRNGSamples=response.nRNGSamples)
        if response._kind != _abi.VALUE_UNIFORM:
            raise ValueError("this build offers UniformValueResponse only (user GLSL cannot be honoured)")
        self._response = response
        self.setParams(t0=t0, binSize=binSize)
        self._nBins = nBins
        self._normalization = normalization
        self._requestedNorm = 1.0
        self._retrieve = retrieve
        self._reduceBlockSize = reduceBlockSize
        self._updateResponse = updateResponse
        self._nHistMax = 0
        self._reducer = None
        if nBins > HistogramReducer.MAX_SHARED_MEMORY_BINS:
            useSharedMemory = False
        self._useSharedMemory = useSharedMemory
        self._nHistLast = [0, 0]

    @property
    def nBins(self) -> int:
        return self._nBins

    @property
    def normalization(self) -> float | None:
        return self._normalization

    @normalization.setter
    def normalization(self, value: float | None) -> None:
        self._normalization = value
        if self._reducer is not None and value is not None:
            self._reducer.normalization = value

    @property
    def reduceBlockSize(self) -> int:
        return self._reduceBlockSize

    @property
    def response(self) -> ValueResponse:
        return self._response

    @property
    def retrieve(self) -> bool:
        return self._retrieve

    @property
    def useSharedMemory(self) -> bool:
        return self._useSharedMemory

    def result(self, i: int) -> NDArray | None:
        return None if self._reducer is None else self._reducer.result(i)

    def prepare(self, config: TraceConfig) -> None:
        self.response.prepare(config)
        self._requestedNorm = config.normalization
        # the tracers run a persistent grid: at most this many CTAs write a partial histogram
        nHistMax = min(-(config.capacity // -config.blockSize), _maxGridBlocks())
        if self._reducer is None:
            self._nHistMax = max(self._nHistMax, nHistMax)
        elif nHistMax > self._reducer.capacity:
            raise RuntimeError(f"Tracer may produce up to {nHistMax} intermediate histograms, "
                               f"but the response can only fit {self._reducer.capacity}")

    def updateConfig(self, config: TraceConfig) -> None:
        if self._reducer is None:
            self._requestedNorm = config.normalization
        else:
            # unused partials stay zero (they are cleared by every reduce), so always sum all
            self._reducer.nHist = self._reducer.capacity
            if self.normalization is None:
                self._reducer.normalization = config.normalization

    def _initReducer(self) -> HistogramReducer:
        norm = self.normalization if self.normalization is not None else self._requestedNorm
        with _device.useDevice(self._device):
            return HistogramReducer(nBins=self.nBins, nHist=max(self._nHistMax, 1), retrieve=self.retrieve,
                                    normalization=norm, blockSize=self._reduceBlockSize,
                                    useSharedMemory=self.useSharedMemory)

    def _abiResponse(self, i: int, nBlocks: int) -> _abi.Response:
        if self._reducer is None:
            self._reducer = self._initReducer()
        p = self._block("ResponseParams", i)
        r = _abi.Response()
        r.kind = self._kind
        r.value_kind = self.response._kind
        r.hist.t0 = p.t0
        r.hist.binSize = p.binSize
        r.n_bins = self.nBins
        r.hist_capacity = self._reducer.capacity
        r.hist_partials = self._reducer.histIn.data_ptr()
        return r

    def _setRows(self, i: int, rows: int) -> None:
        """Called by the tracer while it assembles launch i: partial histograms it will write."""
        self._reducer._block("Params", i).nHist = rows

    def update(self, i: int) -> None:
        if self._reducer is None:
            self._reducer = self._initReducer()
        super().update(i)
        self._reducer.update(i)
        if self._updateResponse:
            self.response.update(i)

    def run(self, i: int) -> list[Command]:
        if self._reducer is None:
            self._reducer = self._initReducer()
        return self._reducer.run(i)


class KernelHistogramHitResponse(HistogramHitResponse):
    """
    Histogram in which every hit is smeared by a Gaussian kernel of width ``kernelBandwidth``
    over the bins within ``kernelSupport`` of its time: a binned kernel density estimate
    (src/theia/response.py:1424-1673, sh/response.histogram.kernel.glsl). Partial histograms and
    their reduction work as for ``HistogramHitResponse``.
    """

    name = "Kernel Histogram Hit Response"
    _kind = _abi.RESPONSE_KERNEL_HISTOGRAM

    class Params(Structure):
        _fields_ = [("t0", c_float), ("binSize", c_float), ("kernelBandwidth", c_float), ("kernelSupport", c_float)]

    def __init__(self, response: ValueResponse, *, nBins: int = 100, t0: float = 0.0 * u.ns,
                 binSize: float = 1.0 * u.ns, kernelBandwidth: float = 1.0 * u.ns, kernelSupport: float = 3.0 * u.ns,
                 normalization: float | None = None, retrieve: bool = True, reduceBlockSize: int = 128,
                 updateResponse: bool = True, useSharedMemory: bool = True) -> None:
        super().__init__(response, nBins=nBins, t0=t0, binSize=binSize, normalization=normalization, retrieve=retrieve,
                         reduceBlockSize=reduceBlockSize, updateResponse=updateResponse,
                         useSharedMemory=useSharedMemory)
        self.setParams(kernelBandwidth=kernelBandwidth, kernelSupport=kernelSupport)

    def _abiResponse(self, i: int, nBlocks: int) -> _abi.Response:
        r = super()._abiResponse(i, nBlocks)
        p = self._block("ResponseParams", i)
        r.kernel_bandwidth = p.kernelBandwidth
        r.kernel_support = p.kernelSupport
        return r


class StoreTimeHitResponse(HitResponse):
    """
    Keeps a hit with the probability its value response assigns to it and stores its arrival time
    and, optionally, the id of the object it hit (src/theia/response.py:656-797,
    sh/response.time.store.glsl). Meant for single photon tracing, where the value models the
    detection probability. Consumes one random number per hit. ``capacity`` = 0 sizes the queue
    from what the tracer reports. ``result(i)`` is a queue view with ``count``, ``time`` and
    ``objectId``.
    """

    name = "Store Time Hit Response"
    _kind = _abi.RESPONSE_STORE_TIME

    def __init__(self, response: ValueResponse, *, capacity: int = 0, retrieve: bool = True,
                 storeObjectId: bool = True, updateResponse: bool = True) -> None:
        super().__init__(nRNGSamples=response.nRNGSamples + 1)
        if response._kind != _abi.VALUE_UNIFORM:
            raise ValueError("this build offers UniformValueResponse only (user GLSL cannot be honoured)")
        self._response = response
        self._capacity = capacity
        self._retrieve = retrieve
        self._storeObjectId = storeObjectId
        self._updateResponse = updateResponse
        self._queue = IOQueue(HitTimeAndIdItem if storeObjectId else HitTimeItem, mode="retrieve")
        if capacity > 0:
            self._queue.initialize(capacity)
        self._tags = None

    @property
    def queue(self) -> IOQueue:
        """Queue to be filled with the arrival times and optionally the object ids"""
        return self._queue

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def response(self) -> ValueResponse:
        return self._response

    @property
    def retrieve(self) -> bool:
        return self._retrieve

    @property
    def storeObjectId(self) -> bool:
        return self._storeObjectId

    def result(self, i: int):
        """The retrieved i-th queue; None if ``retrieve`` is False or nothing ran yet"""
        if not self.retrieve or not self.queue.initialized:
            return None
        return self.queue.view(i)

    def prepare(self, config: TraceConfig) -> None:
        self.response.prepare(config)
        maxHits = config.maxHitsPerThread * config.capacity
        if self.queue.initialized and maxHits > self.capacity:
            warn(f"The provided queue's capacity ({self.capacity}) is less than the max hit count reported by "
                 f"the tracer ({maxHits}). Hits beyond the capacity are counted but not stored.")
        elif not self.queue.initialized:
            self._capacity = max(self._capacity, maxHits)

    def enableTags(self) -> None:
        """B200 extension (test seam): records which path and event produced each stored time."""
        self._wantTags = True

    def tags(self) -> tuple[NDArray, NDArray]:
        return self._tags[0].cpu().numpy(), self._tags[1].cpu().numpy()

    def _ensure(self) -> None:
        if not self.queue.initialized:
            if self.capacity <= 0:
                raise RuntimeError(f"Cannot create a queue of size {self.capacity}")
            self.queue.initialize(self.capacity)
        if getattr(self, "_wantTags", False) and self._tags is None:
            self._tags = [torch.zeros(self.capacity, dtype=torch.int32, device=self._device) for _ in range(2)]

    def _abiResponse(self, i: int, nBlocks: int) -> _abi.Response:
        self._ensure()
        r = _abi.Response()
        r.kind = _abi.RESPONSE_STORE_TIME
        r.value_kind = self.response._kind
        r.value_queue = self.queue.tensor.data_ptr()
        r.value_capacity = self.capacity
        r.store_object_id = int(self._storeObjectId)
        if self._tags is not None:
            r.tag_path = self._tags[0].data_ptr()
            r.tag_event = self._tags[1].data_ptr()
        return r

    def update(self, i: int) -> None:
        super().update(i)
        if self._updateResponse:
            self.response.update(i)

    def run(self, i: int) -> list[Command]:
        self._ensure()
        return self.queue.run(i) if self.retrieve else []


class ObjectHistogramHitResponse(HitResponse):
    """
    B200 extension: one time histogram per detector object, ``result(i)[objectId, bin]`` - the
    per module light curves of a detector array (config C5), which the reference can only
    obtain from ``HitRecorder`` plus binning on the host. Hits whose ``objectId`` lies outside
    ``[0, nObjects)`` or whose time lies outside the histogram are dropped. Binning and value are
    those of ``HistogramHitResponse`` (sh/response.histogram.glsl:59-68, sh/response.uniform.glsl);
    lanes of a warp that hit the same (object, bin) are combined and issue one float atomic on
    the ``[nObjects, nBins]`` array in HBM.
    """

    name = "Object Histogram Hit Response"
    _kind = _abi.RESPONSE_OBJECT_HISTOGRAM

    class Params(Structure):
        _fields_ = [("t0", c_float), ("binSize", c_float)]

    def __init__(self, response: ValueResponse, *, nObjects: int, nBins: int = 100, t0: float = 0.0 * u.ns,
                 binSize: float = 1.0 * u.ns, normalization: float | None = None, retrieve: bool = True,
                 updateResponse: bool = True) -> None:
        super().__init__({"ResponseParams": self.Params}, {"normalization"}, nRNGSamples=response.nRNGSamples)
        if response._kind != _abi.VALUE_UNIFORM:
            raise ValueError("this build offers UniformValueResponse only (user GLSL cannot be honoured)")
        if nObjects <= 0 or nBins <= 0:
            raise ValueError("nObjects and nBins must be positive")
        self._response = response
        self._nObjects = nObjects
        self._nBins = nBins
        self.normalization = normalization
        self._requestedNorm = 1.0
        self._retrieve = retrieve
        self._updateResponse = updateResponse
        self.setParams(t0=t0, binSize=binSize)
        n = nObjects * nBins
        self._hist = _device.alloc(4 * n, self._device).view(torch.float32)[:n]
        self._buffers = [_device.allocPinned(4 * n).view(torch.float32)[:n] for _ in range(2)] if retrieve else None

    @property
    def nObjects(self) -> int:
        return self._nObjects

    @property
    def nBins(self) -> int:
        return self._nBins

    @property
    def response(self) -> ValueResponse:
        return self._response

    @property
    def retrieve(self) -> bool:
        return self._retrieve

    @property
    def tensor(self) -> torch.Tensor:
        """Device array ``[nObjects * nBins]`` the tracer accumulates into (un-normalised)"""
        return self._hist

    def result(self, i: int) -> NDArray | None:
        if self._buffers is None:
            return None
        return self._buffers[i].numpy().reshape(self._nObjects, self._nBins)

    def prepare(self, config: TraceConfig) -> None:
        self.response.prepare(config)
        self._requestedNorm = config.normalization

    def updateConfig(self, config: TraceConfig) -> None:
        self._requestedNorm = config.normalization

    def _abiResponse(self, i: int, nBlocks: int) -> _abi.Response:
        p = self._block("ResponseParams", i)
        r = _abi.Response()
        r.kind = _abi.RESPONSE_OBJECT_HISTOGRAM
        r.value_kind = self.response._kind
        r.hist.t0 = p.t0
        r.hist.binSize = p.binSize
        r.n_bins = self._nBins
        r.object_hist = self._hist.data_ptr()
        r.n_objects = self._nObjects
        return r

    def update(self, i: int) -> None:
        super().update(i)
        if self._updateResponse:
            self.response.update(i)

    def run(self, i: int) -> list[Command]:
        def cmd() -> None:
            norm = self.normalization if self.normalization is not None else self._requestedNorm
            if self._buffers is not None:
                if norm != 1.0:
                    self._hist.mul_(norm)
                self._buffers[i].copy_(self._hist, non_blocking=True)
            self._hist.zero_()
        return [cmd]


def _maxGridBlocks() -> int:
    """Upper bound of CTAs a tracer launch uses (persistent grid, see csrc/trace.cu)."""
    return 148 * 16


class HitReplay(PipelineStage):
    """
    Feeds hits from a host queue to a response: the seam that tests a response without a
    tracer (src/theia/response.py:278-422, sh/response.replay.glsl).
    """

    name = "Hit Replay"

    def __init__(self, capacity: int, response: HitResponse, *, rng: RNG | None = None, normalization: float = 1.0,
                 polarized: bool = False, blockSize: int = 128, code: bytes | None = None) -> None:
        if response.nRNGSamples > 0 and rng is None:
            raise ValueError("response requires a rng but none was given!")
        if isinstance(response, HitRecorder) and response.polarized != polarized:
            raise RuntimeError("mismatch in hit queue definition")
        super().__init__()
        self._blockSize = blockSize
        self._capacity = capacity
        self._polarized = polarized
        self._response = response
        self._rng = rng
        self._config = TraceConfig(capacity, blockSize, capacity, 1, normalization, polarized)
        response.prepare(self._config)
        self._queue = IOQueue(PolarizedHitItem if polarized else HitItem, capacity, mode="update")

    @property
    def blockSize(self) -> int:
        return self._blockSize

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def polarized(self) -> bool:
        return self._polarized

    @property
    def queue(self) -> IOQueue:
        return self._queue

    @property
    def rng(self) -> RNG | None:
        return self._rng

    @property
    def response(self) -> HitResponse:
        return self._response

    def collectStages(self) -> list[PipelineStage]:
        if self.rng is not None:
            return [self.rng, self, self.response]
        return [self, self.response]

    def _finishParams(self, i: int) -> None:
        super()._finishParams(i)
        self.response.updateConfig(self._config)

    def run(self, i: int) -> list[Command]:
        def cmd() -> None:
            if self._device != "cuda":
                raise RuntimeError("HitReplay needs a CUDA device (no CPU fallback)")
            nBlocks = -(self.capacity // -self.blockSize)
            r = self.response._abiResponse(i, nBlocks)
            rng = self.rng._abiParams(i) if self.rng is not None else None
            _lib.check(_lib.load().theia_hit_replay_rng(
                self.queue.tensor.data_ptr(), self.capacity, int(self.polarized), byref(r),
                ctypes.addressof(rng) if rng is not None else None, self.blockSize, _device.currentStream("cuda")))

        return [*self.queue.run(i), cmd]
