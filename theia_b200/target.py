"""
Targets of the volume tracer and target guides of the scene tracer (API of the reference's
src/theia/target.py; sphere shapes).
"""

from __future__ import annotations

from ctypes import Structure, c_float

import numpy as np

import theia_b200.units as u
from theia_b200 import _abi
from theia_b200.glsl import vec3
from theia_b200.pipeline import SourceCodeMixin

__all__ = ["InnerSphereTarget", "SphereTarget", "SphereTargetGuide", "Target", "TargetGuide"]


class Target(SourceCodeMixin):
    """
    Base class of targets: the detector stand-in of the volume tracer. Device side it offers
    sampleTarget, intersectTarget and isOccludedByTarget (src/theia/target.py:36-75).
    """

    name = "Target"

    def __init__(self, *, nRNGSamples: int, params: dict[str, type[Structure]] = {}, extra: set[str] = set()) -> None:
        super().__init__(params, extra)
        self._nRNGSamples = nRNGSamples

    @property
    def nRNGSamples(self) -> int:
        return self._nRNGSamples


class _SphereParams(Structure):
    _fields_ = [("position", vec3), ("radius", c_float), ("_invPos", vec3), ("_invRad", c_float), ("_prob", c_float)]


class _SphereTargetBase(Target):
    _areaFactor = 2.0

    def __init__(self, *, position=(0.0, 0.0, 0.0), radius: float = 1.0 * u.m) -> None:
        super().__init__(nRNGSamples=2, params={"TargetParams": self.TargetParams})
        self.setParams(position=position, radius=radius)

    def _finishParams(self, i: int) -> None:
        r = self.getParam("radius")
        x, y, z = self.position
        # sample probability per area and the world -> unit sphere transformation
        self.setParams(_prob=1.0 / (self._areaFactor * np.pi * r**2), _invPos=(-x / r, -y / r, -z / r), _invRad=1.0 / r)


class SphereTarget(_SphereTargetBase):
    """
    Sphere seen from outside; samples the part visible from the observer. Hits are reported
    on the unit sphere in object space (src/theia/target.py:78-139, sh/target.sphere.glsl).
    """

    name = "Sphere Target"
    _kind = _abi.TARGET_SPHERE
    _areaFactor = 2.0  # hemisphere: 1 / (2 pi r^2)

    class TargetParams(_SphereParams):
        pass


class InnerSphereTarget(_SphereTargetBase):
    """
    Sphere seen from inside, sampled uniformly (src/theia/target.py:142-199,
    sh/target.sphere.inner.glsl).
    """

    name = "Inner Sphere Target"
    _kind = _abi.TARGET_INNER_SPHERE
    _areaFactor = 4.0  # full sphere: 1 / (4 pi r^2)

    class TargetParams(_SphereParams):
        pass


class TargetGuide(SourceCodeMixin):
    """
    Base class of target guides: direction samplers used by the scene tracer for next event
    estimation (src/theia/target.py:427-467).
    """

    name = "Target Guide"

    def __init__(self, *, nRNGSamples: int, params: dict[str, type[Structure]] = {}, extra: set[str] = set()) -> None:
        super().__init__(params, extra)
        self._nRNGSamples = nRNGSamples

    @property
    def nRNGSamples(self) -> int:
        return self._nRNGSamples


class SphereTargetGuide(TargetGuide):
    """
    Samples directions inside the cone subtended by a sphere; rays are traced through it up
    to its far side (src/theia/target.py:470-525, sh/target_guide.sphere.glsl).
    """

    name = "Sphere Target Guide"
    _kind = _abi.GUIDE_SPHERE

    class TargetGuideParams(Structure):
        _fields_ = [("position", vec3), ("radius", c_float), ("_prob", c_float)]

    def __init__(self, *, position=(0.0, 0.0, 0.0), radius: float = 1.0 * u.m) -> None:
        super().__init__(nRNGSamples=2, params={"TargetGuideParams": self.TargetGuideParams})
        self.setParams(position=position, radius=radius)

    def _finishParams(self, i: int) -> None:
        r = self.getParam("radius")
        self.setParam("_prob", 1.0 / (2.0 * np.pi * r**2))
