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
    "HitResponse", "PolarizedHitItem", "TraceConfig", "UniformValueResponse", "ValueItem", "ValueResponse",
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
        super().__init__({"ResponseParams": self.Params}, {"normalization"}, nRNGSamples=response.nRNGSamples)
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
        r.kind = _abi.RESPONSE_HISTOGRAM
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
            _lib.check(_lib.load().theia_hit_replay(
                self.queue.tensor.data_ptr(), self.capacity, int(self.polarized), byref(r), self.blockSize,
                _device.currentStream("cuda")))

        return [*self.queue.run(i), cmd]
