"""
Random number generators (API of the reference's src/theia/random.py).

``PhiloxRNG`` reproduces the reference counter for counter: draw ``i`` of stream ``s`` is
``Philox4x32-10(ctr = offset + 4 i, key = key + s)[i mod 4]`` mapped to [0, 1)
(sh/random.philox.glsl:15-77, see DESIGN.md "Philox").
"""

from __future__ import annotations

import warnings
from ctypes import Structure, byref, c_uint32
from os import urandom

import numpy as np
import torch

from theia_b200 import _abi, _device, _lib
from theia_b200.pipeline import Command, PipelineStage, SourceCodeMixin

__all__ = ["Counter", "Key", "PhiloxRNG", "RNG", "RNGBufferSink"]


class RNG(SourceCodeMixin):
    """Base class of random number generators used by other stages."""

    name = "RNG"

    def __init__(self, params: dict[str, type[Structure]] = {}, extras: set[str] = set()) -> None:
        super().__init__(params, extras)


class Key(Structure):
    """64 bit key as two little endian words (src/theia/random.py:202-212)"""

    _fields_ = [("lo", c_uint32), ("hi", c_uint32)]

    @property
    def value(self) -> int:
        return self.lo + (self.hi << 32)

    @value.setter
    def value(self, value: int) -> None:
        self.lo = value & 0xFFFFFFFF
        self.hi = value >> 32 & 0xFFFFFFFF


class Counter(Structure):
    """128 bit counter as four little endian words (src/theia/random.py:215-225)"""

    _fields_ = [("word", c_uint32 * 4)]

    @property
    def value(self) -> int:
        return sum(self.word[i] << (32 * i) for i in range(4))

    @value.setter
    def value(self, value: int) -> None:
        for i in range(4):
            self.word[i] = value >> (32 * i) & 0xFFFFFFFF


class PhiloxRNG(RNG):
    """
    Philox 4x32-10 generator (src/theia/random.py:228-282).

    Parameters
    ----------
    key: int | None
        Base key; stream ``s`` uses ``key + s``. A random key is drawn (with a warning) if None.
    offset: int
        Offset added to the 128 bit counter of every stream.
    autoAdvance: int
        Amount added to ``offset`` on every ``update``; set it to ``tracer.nRNGSamples`` so
        consecutive batches use fresh numbers.
    """

    name = "Philox RNG"

    class PhiloxParams(Structure):
        _fields_ = [("key", Key), ("offset", Counter)]

    def __init__(self, *, key: int | None = None, offset: int = 0, autoAdvance: int = 0) -> None:
        super().__init__({"PhiloxParams": self.PhiloxParams}, {"autoAdvance"})
        if key is None:
            key = int.from_bytes(urandom(8), "little")
            warnings.warn(f"Random RNG key generated: 0x{key:08X}")
        self.setParams(key=key, offset=offset, autoAdvance=autoAdvance)

    @property
    def autoAdvance(self) -> int:
        return self._autoAdvance

    @autoAdvance.setter
    def autoAdvance(self, value: int) -> None:
        self._autoAdvance = value

    def _finishParams(self, i: int) -> None:
        if self.autoAdvance != 0:
            self.setParam("offset", self.getParam("offset") + self.autoAdvance)
        super()._finishParams(i)

    def _abiParams(self, i: int) -> _abi.PhiloxParams:
        """The committed parameters of configuration i in C-ABI form"""
        block = self._block("PhiloxParams", i)
        p = _abi.PhiloxParams()
        p.key[0], p.key[1] = block.key.lo, block.key.hi
        for k in range(4):
            p.offset[k] = block.offset.word[k]
        return p


class RNGBufferSink(PipelineStage):
    """
    Fills a tensor with ``streams x samples`` random numbers (src/theia/random.py:44-199):
    element ``[s, c]`` is draw ``baseCount + c`` of stream ``baseStream + s``.
    """

    name = "RNG Sink"

    class Params(Structure):
        _fields_ = [("baseStream", c_uint32), ("baseCount", c_uint32), ("_streams", c_uint32), ("_samples", c_uint32)]

    def __init__(self, generator: RNG, streams: int = 256, samples: int = 256, *, baseStream: int = 0,
                 baseCount: int = 0, sampleDim: int = 1, blockSize: tuple[int, int, int] = (32, 4, 16),
                 code: bytes | None = None) -> None:
        super().__init__({"Params": self.Params})
        if not 1 <= sampleDim <= 2:
            raise ValueError("Unsupported sample dimension requested!")
        if not isinstance(generator, PhiloxRNG):
            raise ValueError("only PhiloxRNG is available in this build")
        self._capacity = streams * samples * sampleDim
        self._generator = generator
        self._blockSize = blockSize
        self._shape = (streams, samples, sampleDim)
        self.setParams(baseStream=baseStream, baseCount=baseCount, _streams=streams, _samples=samples)
        self._tensor = torch.zeros(self._capacity, dtype=torch.float32, device=self._device)

    @property
    def blockSize(self):
        return self._blockSize

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def generator(self) -> RNG:
        return self._generator

    @property
    def shape(self):
        return self._shape

    @property
    def streams(self) -> int:
        return self.getParam("_streams")

    @property
    def samples(self) -> int:
        return self.getParam("_samples")

    @property
    def tensor(self) -> torch.Tensor:
        return self._tensor

    def run(self, i: int) -> list[Command]:
        def cmd() -> None:
            if self._device != "cuda":
                raise RuntimeError("RNGBufferSink needs a CUDA device (no CPU fallback)")
            p = self._block("Params", i)
            rng = self._generator._abiParams(i)
            _lib.check(_lib.load().theia_philox_fill(
                byref(rng), p.baseStream, p.baseCount, p._streams, p._samples, self._shape[2],
                self._tensor.data_ptr(), _device.currentStream("cuda")))

        return [cmd]
