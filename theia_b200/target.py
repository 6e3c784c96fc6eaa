"""
Targets of the volume tracer and target guides of the scene tracer (API of the reference's
src/theia/target.py: sphere, inner sphere, flat and disk shapes).
"""

from __future__ import annotations

from ctypes import Structure, c_float

import numpy as np

import theia_b200.units as u
from theia_b200 import _abi
from theia_b200.glsl import mat3, vec3
from theia_b200.pipeline import SourceCodeMixin

__all__ = [
    "DiskTarget",
    "DiskTargetGuide",
    "FlatTarget",
    "FlatTargetGuide",
    "InnerSphereTarget",
    "SphereTarget",
    "SphereTargetGuide",
    "Target",
    "TargetGuide",
]


def _viewMatrix(direction, up):
    """inner 3x3 of the reference's Transform.View (src/theia/scene.py:190-231): columns x, y, z"""
    from theia_b200.scene import Transform

    if up is None:
        up = (0.0, 1.0, 0.0)
    return Transform.View(direction=direction, up=up).innerMatrix


class _FlatParams(Structure):
    # src/theia/target.py:262-270 / 574-582 ("length" is called "height" by the guide)
    _fields_ = [("width", c_float), ("length", c_float), ("position", vec3), ("_normal", vec3), ("_prob", c_float),
                ("_objToWorld", mat3)]


class _FlatGuideParams(Structure):
    _fields_ = [("width", c_float), ("height", c_float), ("position", vec3), ("_normal", vec3), ("_prob", c_float),
                ("_objToWorld", mat3)]


class _DiskParams(Structure):
    # src/theia/target.py:363-370 / 672-679
    _fields_ = [("position", vec3), ("radius", c_float), ("_normal", vec3), ("_prob", c_float), ("_objToWorld", mat3)]


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


class FlatTarget(Target):
    """
    Rectangle in the local xy plane (width along x, length along y), normal along local z;
    ``direction`` is the world space normal, ``up`` fixes the local y axis
    (src/theia/target.py:202-306, sh/target.flat.glsl). Both sides detect; never occludes.
    """

    name = "Flat Target"
    _kind = _abi.TARGET_FLAT

    class TargetParams(_FlatParams):
        pass

    def __init__(self, *, width: float = 1.0 * u.cm, length: float = 1.0 * u.cm, position=(0.0, 0.0, 0.0),
                 direction=(0.0, 0.0, 1.0), up=(0.0, 1.0, 0.0)) -> None:
        super().__init__(nRNGSamples=2, params={"TargetParams": self.TargetParams}, extra={"direction", "up"})
        self.direction = direction
        self.up = up
        self.setParams(width=width, length=length, position=position)

    def _finishParams(self, i: int) -> None:
        m = _viewMatrix(self.direction, self.up)
        self.setParam("_objToWorld", m.T)  # column major
        self.setParam("_normal", m[:, 2])
        self.setParam("_prob", 1.0 / (self.getParam("width") * self.getParam("length")))


class DiskTarget(Target):
    """
    Disk in the local xy plane, normal along local z (src/theia/target.py:309-424,
    sh/target.disk.glsl). Both sides detect; never occludes.
    """

    name = "Disk Target"
    _kind = _abi.TARGET_DISK

    class TargetParams(_DiskParams):
        pass

    def __init__(self, *, position=(0.0, 0.0, 0.0), radius: float = 1.0 * u.m, normal=(0.0, 0.0, 1.0), up=None) -> None:
        super().__init__(nRNGSamples=2, params={"TargetParams": self.TargetParams}, extra={"normal", "up"})
        self.normal = normal
        self.up = up
        self.setParams(position=position, radius=radius)

    def _finishParams(self, i: int) -> None:
        m = _viewMatrix(self.normal, self.up)
        self.setParam("_objToWorld", m.T)  # column major
        self.setParam("_normal", m[:, 2])  # normalised
        self.setParam("_prob", 1.0 / (np.pi * self.getParam("radius") ** 2))


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


class FlatTargetGuide(TargetGuide):
    """
    Samples points on a rectangle; only rays arriving against its normal count
    (src/theia/target.py:528-636, sh/target_guide.flat.glsl).
    """

    name = "Flat Target Guide"
    _kind = _abi.GUIDE_FLAT

    class TargetGuideParams(_FlatGuideParams):
        pass

    def __init__(self, *, width: float = 1.0 * u.m, height: float = 1.0 * u.m, position=(0.0, 0.0, 0.0),
                 normal=(0.0, 0.0, 1.0), up=(0.0, 1.0, 0.0)) -> None:
        super().__init__(nRNGSamples=2, params={"TargetGuideParams": self.TargetGuideParams}, extra={"normal", "up"})
        self.normal = normal
        self.up = up
        self.setParams(width=width, height=height, position=position)

    def _finishParams(self, i: int) -> None:
        m = _viewMatrix(self.normal, self.up)
        self.setParam("_objToWorld", m.T)  # column major
        self.setParam("_normal", m[:, 2])
        self.setParam("_prob", 1.0 / (self.getParam("width") * self.getParam("height")))


class DiskTargetGuide(TargetGuide):
    """
    Samples points on a disk; only rays arriving against its normal count
    (src/theia/target.py:639-735, sh/target_guide.disk.glsl).
    """

    name = "Disk Target Guide"
    _kind = _abi.GUIDE_DISK

    class TargetGuideParams(_DiskParams):
        pass

    def __init__(self, *, position=(0.0, 0.0, 0.0), radius: float = 1.0 * u.m, normal=(0.0, 0.0, 1.0)) -> None:
        super().__init__(nRNGSamples=2, params={"TargetGuideParams": self.TargetGuideParams}, extra={"normal"})
        self.normal = normal
        self.setParams(position=position, radius=radius)

    def _finishParams(self, i: int) -> None:
        r = self.getParam("radius")
        self.setParam("_prob", 1.0 / (np.pi * r**2))
        z = np.array(self.normal, dtype=np.float64)
        z = z / np.sqrt(np.square(z).sum(-1))
        self.setParam("_normal", z)
        # frame from the normal (PBRT 4th ed., ch. 3.3.3), exactly as the reference builds it
        # (src/theia/target.py:716-734). NB the reference stores this matrix WITHOUT the
        # transposition its other shapes apply, so the shader reads its transpose; kept as is,
        # because the sampled positions of the reference are what has to be reproduced.
        vx, vy, vz = z
        sign = np.copysign(1.0, vz)
        a = -1.0 / (sign + vz)
        b = vx * vy * a
        x = np.array([1.0 + sign * vx**2 * a, sign * b, -sign * vx])
        y = np.array([b, sign + vy**2 * a, -vy])
        self.setParam("_objToWorld", np.stack([x, y, z], -1))
