"""Helpers shared by tests and examples (reference: src/theia/testing.py:641-656)."""

from __future__ import annotations

import numpy as _np

from theia_b200 import _abi
from theia_b200.pipeline import PipelineStage
from theia_b200.material import (
    HenyeyGreensteinPhaseFunction,
    KokhanovskyOceanWaterPhaseMatrix,
    MediumModel,
    WaterBaseModel,
)

__all__ = ["TargetGuideSampler", "TargetSampler", "WaterTestModel"]


class WaterTestModel(WaterBaseModel, HenyeyGreensteinPhaseFunction, KokhanovskyOceanWaterPhaseMatrix, MediumModel):
    """Sea water at 5 C, 1000 dbar, 35 PSU with HG g = 0.6 and Kokhanovsky's phase matrix (Voss fit)."""

    ModelName = "water"

    def __init__(self) -> None:
        WaterBaseModel.__init__(self, 5.0, 1000.0, 35.0)
        HenyeyGreensteinPhaseFunction.__init__(self, 0.6)
        KokhanovskyOceanWaterPhaseMatrix.__init__(self, p90=0.66, theta0=0.25, alpha=4.0, xi=25.6)


# ------------------------------------------------------------------ sampler stages (src/theia/testing.py:429-638)
def _rows(buffer, fields, n):
    """splits the float32 rows of a unit evaluation into the named fields of the reference's Item structure"""
    import numpy as np
    data = buffer.numpy().view(np.float32).reshape(n, -1)
    out, at = {}, 0
    for name, width, dtype in fields:
        col = data[:, at:at + width]
        if dtype is not np.float32:
            col = col.view(dtype)
        out[name] = col[:, 0] if width == 1 else col
        at += width
    return out


class _UnitSampler(PipelineStage):
    """Common part of TargetSampler / TargetGuideSampler: one theia_unit_eval launch + retrieval"""

    _unit = 0
    _itemFields: list = []

    def __init__(self, capacity: int, component, *, rng, sampleBox) -> None:
        import torch
        from theia_b200 import _device, _units
        super().__init__()
        self._component = component
        self._rng = rng
        self._batch = capacity
        self._box = (*sampleBox.lowerCorner, *sampleBox.upperCorner)
        self._width = _units.WIDTHS[self._unit][1]
        self._tensor = torch.zeros(capacity * self._width, dtype=torch.float32, device=self._device)
        self._buffer = [_device.allocPinned(4 * capacity * self._width) for _ in range(2)]

    @property
    def batchSize(self) -> int:
        """Number of samples to draw per batch"""
        return self._batch

    @property
    def rng(self):
        """Generator for creating random numbers"""
        return self._rng

    def collectStages(self) -> list[PipelineStage]:
        return [self._rng, self._component, self]

    def getResults(self, i: int) -> dict:
        return _rows(self._buffer[i][: 4 * self._batch * self._width], self._itemFields, self._batch)

    def _unitArgs(self, i: int):
        from theia_b200 import _units
        kw = {"target": self._component} if self._unit == _abi.UNIT_TARGET_SAMPLE else {"guide": self._component}
        A = _units.unitArgs(self._unit, self._batch, rng=self._rng, i=i, params=self._box, **kw)
        A.out = self._tensor.data_ptr()
        return A

    def run(self, i: int):
        from ctypes import byref
        from theia_b200 import _device, _lib

        def cmd() -> None:
            if self._device != "cuda":
                raise RuntimeError(f"{type(self).__name__} needs a CUDA device (no CPU fallback)")
            A = self._unitArgs(i)
            A.stream = _device.currentStream("cuda")
            _lib.check(_lib.load().theia_unit_eval(byref(A)))
            self._buffer[i][: 4 * self._batch * self._width].view(self._tensor.dtype).copy_(self._tensor, non_blocking=True)

        return [cmd]


class TargetSampler(_UnitSampler):
    """
    Test program for sampling targets (src/theia/testing.py:537-638, sh/target.sample.glsl): per sample an
    observer inside ``sampleBox`` and a direction are drawn, then ``sampleTarget``, ``intersectTarget`` and
    ``isOccludedByTarget`` are evaluated and checked for self consistency (``sampleError`` / ``hitError``).
    """

    name = "Target Sampler"
    _unit = _abi.UNIT_TARGET_SAMPLE
    _itemFields = [("observer", 3, _np.float32), ("direction", 3, _np.float32), ("samplePos", 3, _np.float32),
               ("sampleNrm", 3, _np.float32), ("sampleProb", 1, _np.float32), ("sampleValid", 1, _np.uint32),
               ("sampleError", 1, _np.int32), ("hitPos", 3, _np.float32), ("hitNrm", 3, _np.float32),
               ("hitProb", 1, _np.float32), ("hitValid", 1, _np.uint32), ("hitError", 1, _np.int32),
               ("occluded", 1, _np.uint32)]

    def __init__(self, capacity: int, target, *, rng, sampleBox=None) -> None:
        from theia_b200.scene import RectBBox
        super().__init__(capacity, target, rng=rng, sampleBox=sampleBox or RectBBox((-20.0,) * 3, (20.0,) * 3))

    @property
    def target(self):
        """Target to be sampled"""
        return self._component


class TargetGuideSampler(_UnitSampler):
    """Test program for sampling target guides (src/theia/testing.py:429-534, sh/target_guide.sample.glsl)."""

    name = "Target Guide Sampler"
    _unit = _abi.UNIT_GUIDE_SAMPLE
    _itemFields = [("observer", 3, _np.float32), ("sampleDir", 3, _np.float32), ("sampleDist", 1, _np.float32),
               ("sampleProb", 1, _np.float32), ("evalDir", 3, _np.float32), ("evalDist", 1, _np.float32),
               ("evalProb", 1, _np.float32)]

    def __init__(self, capacity: int, guide, *, rng, sampleBox=None) -> None:
        from theia_b200.scene import RectBBox
        super().__init__(capacity, guide, rng=rng, sampleBox=sampleBox or RectBBox((-20.0,) * 3, (20.0,) * 3))

    @property
    def guide(self):
        """Target guide to be sampled"""
        return self._component
