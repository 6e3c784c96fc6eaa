"""
Unit annotations (API of the reference's src/theia/units.py:60-231).

Internal units: metre, nanosecond, nanometre (wavelength), GeV, radian. ``value * unit``
converts an annotated value to internal units, ``value / unit`` converts back. Tuples are
converted element-wise so that ``(0.0, 100.0) * u.ns`` works.
"""

from __future__ import annotations

from math import pi
from typing import Literal

__all__ = [
    "convert", "km", "m", "cm", "mm", "inch", "ft", "yd", "mi", "per_km", "per_m", "per_cm",
    "per_mm", "c", "s", "ms", "us", "ns", "ps", "um", "nm", "rad", "deg", "eV", "keV", "GeV",
    "TeV", "PeV", "Unit", "InverseUnit", "CompoundUnit",
]

Dimension = Literal["length", "time", "wavelength", "compound", "energy", "angle"]


def _each(value, fn):
    return tuple(fn(v) for v in value) if type(value) is tuple else fn(value)


class Unit:
    """Conversion factor from a named unit to the internal unit of its dimension."""

    # makes numpy hand whole arrays to __rmul__/__rtruediv__
    __array_priority__ = 9000

    def __init__(self, dimension: Dimension, scale: float = 1.0) -> None:
        self._dim = dimension
        self._scale = scale

    @property
    def dimension(self) -> Dimension:
        return self._dim

    @property
    def scale(self) -> float:
        return self._scale

    def __rmul__(self, value):
        f = self.scale
        return _each(value, lambda v: v * f)

    def __rtruediv__(self, value):
        f = self.scale
        return _each(value, lambda v: v / f)


class InverseUnit(Unit):
    """1 / unit, e.g. per metre."""

    def __init__(self, unit: Unit) -> None:
        super().__init__(unit.dimension)
        self._unit = unit

    @property
    def scale(self) -> float:
        return 1.0 / self._unit.scale

    def __rmul__(self, value):
        return value / self._unit

    def __rtruediv__(self, value):
        return value * self._unit


class CompoundUnit(Unit):
    """const * prod(nom) / prod(denom)"""

    def __init__(self, const: float, nom, denom) -> None:
        super().__init__("compound")
        self._const = const
        self._nom = list(nom)
        self._denom = list(denom)

    @property
    def scale(self) -> float:
        f = self._const
        for n in self._nom:
            f = f * n
        for d in self._denom:
            f = f / d
        return f


km = Unit("length", 1000.0)
m = Unit("length", 1.0)
cm = Unit("length", 0.01)
mm = Unit("length", 0.001)
inch = Unit("length", 0.0254)
ft = Unit("length", 0.3048)
yd = Unit("length", 0.9144)
mi = Unit("length", 1609.344)
per_km = InverseUnit(km)
per_m = InverseUnit(m)
per_cm = InverseUnit(cm)
per_mm = InverseUnit(mm)
s = Unit("time", 1.0e9)
ms = Unit("time", 1.0e6)
us = Unit("time", 1.0e3)
ns = Unit("time", 1.0)
ps = Unit("time", 1.0e-3)
c = CompoundUnit(299792458.0, [m], [s])
um = Unit("wavelength", 1000.0)
nm = Unit("wavelength", 1.0)
rad = Unit("angle", 1.0)
deg = Unit("angle", pi / 180.0)
eV = Unit("energy", 1e-6)
keV = Unit("energy", 1e-3)
GeV = Unit("energy", 1.0)
TeV = Unit("energy", 1e3)
PeV = Unit("energy", 1e6)


def convert(data, unit: Unit):
    """Converts data given in internal units to the given unit (``data / unit``)."""
    return data / unit
