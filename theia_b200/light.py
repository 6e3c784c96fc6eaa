"""
Light and wavelength sources (API of the reference's src/theia/light.py, forward path).

Each class owns a parameter block that is byte-identical to the reference's
``ctypes.Structure`` and to the matching ``theia_light_*`` / ``theia_wavelength_*`` struct of
include/theia_b200.h; the device code is selected by ``_kind`` / ``_flags``.
"""

from __future__ import annotations

from ctypes import Structure, c_float

import numpy as np

import theia_b200.units as u
from theia_b200 import _abi
from theia_b200.glsl import vec2, vec3, vec4
from theia_b200.pipeline import SourceCodeMixin

__all__ = [
    "ConeLightSource", "ConstWavelengthSource", "LightSampleItem", "LightSource", "MuonTrackLightSource",
    "ParticleCascadeLightSource", "PencilLightSource", "PolarizedLightSampleItem", "SphericalLightSource",
    "UniformWavelengthSource", "WavelengthSampleItem", "WavelengthSource", "frankTamm",
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
