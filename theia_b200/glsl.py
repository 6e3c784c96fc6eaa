"""
ctypes mirrors of the GLSL vector types used in parameter blocks (scalar layout: 4 byte
alignment, no padding). Stand-in for ``hephaistos.glsl`` (reference: imports in
src/theia/trace.py:5, src/theia/light.py:13, src/theia/target.py:12).
"""

from __future__ import annotations

from ctypes import Structure, c_float, c_uint32

__all__ = ["buffer_reference", "mat3", "uvec2", "uvec4", "vec2", "vec3", "vec4"]


class _Vec(Structure):
    """Base for small vectors: indexable, iterable and assignable from sequences."""

    def __len__(self) -> int:
        return len(self._fields_)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return tuple(self)[i]
        return getattr(self, self._fields_[i][0])

    def __setitem__(self, i, v) -> None:
        setattr(self, self._fields_[i][0], v)

    def __iter__(self):
        return (getattr(self, name) for name, _ in self._fields_)

    def assign(self, value) -> None:
        values = tuple(value)
        if len(values) != len(self._fields_):
            raise ValueError(f"expected {len(self._fields_)} values, got {len(values)}")
        for (name, _), v in zip(self._fields_, values):
            setattr(self, name, v)

    def __repr__(self) -> str:
        return f"{type(self).__name__}{tuple(self)}"


class vec2(_Vec):
    _fields_ = [("x", c_float), ("y", c_float)]


class vec3(_Vec):
    _fields_ = [("x", c_float), ("y", c_float), ("z", c_float)]


class vec4(_Vec):
    _fields_ = [("x", c_float), ("y", c_float), ("z", c_float), ("w", c_float)]


class uvec2(_Vec):
    _fields_ = [("x", c_uint32), ("y", c_uint32)]


class uvec4(_Vec):
    _fields_ = [("x", c_uint32), ("y", c_uint32), ("z", c_uint32), ("w", c_uint32)]


class mat3(Structure):
    """Column major 3x3 matrix (assign the transpose of a numpy matrix)."""

    _fields_ = [("m", c_float * 9)]


class buffer_reference(Structure):
    """
    64 bit device address stored as (lo, hi) words: 4 byte aligned like GLSL's uvec2
    (sh/ray.glsl:27; src/theia/util.py:133-140).
    """

    _fields_ = [("lo", c_uint32), ("hi", c_uint32)]

    @property
    def value(self) -> int:
        return self.lo + (self.hi << 32)

    @value.setter
    def value(self, v: int) -> None:
        v = int(v)
        self.lo = v & 0xFFFFFFFF
        self.hi = (v >> 32) & 0xFFFFFFFF
