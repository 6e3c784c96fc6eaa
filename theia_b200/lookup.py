"""
Look-up tables (API of the reference's src/theia/lookup.py).

Device layout of a table (src/theia/lookup.py:48-73, sh/lookup.glsl:4-12): ``int32 n[ndim]``
followed by ``float32 data[...]`` in row-major order, 4 byte aligned. The kernels interpolate
linearly between equidistant samples on [0, 1].
"""

from __future__ import annotations

from ctypes import memmove
from typing import Final, Literal

import numpy as np
import torch
from numpy.typing import ArrayLike, NDArray

from theia_b200 import _device

__all__ = ["Table", "getTableSize", "sampleTable1D", "sampleTable2D", "uploadTables", "evalTable"]


class Table:
    """Values at equidistant sampling points, stored as float32 with a shape header."""

    ALIGNMENT: Final[int] = 4

    def __init__(self, data: ArrayLike) -> None:
        self._data = np.ascontiguousarray(data, dtype=np.float32)
        self._header = np.array(self._data.shape, dtype=np.int32)

    @property
    def data(self) -> NDArray[np.float32]:
        return self._data

    @property
    def nbytes(self) -> int:
        return self._data.nbytes + self._header.nbytes

    def tobytes(self) -> bytes:
        return self._header.tobytes() + self._data.tobytes()

    def copy(self, ptr: int) -> int:
        """Copies header and data to the host address ``ptr``; returns the bytes written."""
        memmove(ptr, self._header.ctypes.data, self._header.nbytes)
        memmove(ptr + self._header.nbytes, self._data.ctypes.data, self._data.nbytes)
        return self.nbytes

    def upload(self) -> torch.Tensor:
        """Uploads the table; returns the buffer holding it."""
        host = np.frombuffer(self.tobytes(), dtype=np.uint8)
        t = _device.alloc(self.nbytes)
        t[: self.nbytes].copy_(torch.from_numpy(host.copy()))
        return t


def getTableSize(a: ArrayLike | tuple[int, ...] | None) -> int:
    """
    Bytes reserved for a table of the given shape (0 for None). Restates the reference's
    arithmetic (src/theia/lookup.py:84-96): ``(ndim + sum(shape)) * 4``, which equals the
    true size for 1D tables, the only kind media use.
    """
    if a is None:
        return 0
    if type(a) is not tuple:
        a = np.shape(a)
    if len(a) == 0:
        raise RuntimeError("table cannot have zero shape!")
    return (len(a) + sum(a)) * 4


def uploadTables(data: list[NDArray]) -> tuple[torch.Tensor, list[int]]:
    """Packs the tables back to back into one buffer; returns it and the table addresses."""
    tables = [Table(d) for d in data]
    blob = b"".join(t.tobytes() for t in tables)
    tensor = _device.alloc(len(blob))
    tensor[: len(blob)].copy_(torch.from_numpy(np.frombuffer(blob, dtype=np.uint8).copy()))
    adr, base = [], tensor.data_ptr()
    for t in tables:
        adr.append(base)
        base += t.nbytes
    return tensor, adr


def _grid(values, boundary, n):
    if boundary is None:
        return np.linspace(values.min(), values.max(), n)
    if type(boundary) is tuple and len(boundary) == 2:
        return np.linspace(boundary[0], boundary[1], n)
    raise RuntimeError("Cant parse given boundaries!")


def sampleTable1D(data, nx: int = 1024, *, boundary=None, mode: Literal["linear", "cubic"] = "linear") -> Table:
    """Resamples (x, f(x)) rows onto nx equidistant points."""
    data = np.asarray(data)
    x = _grid(data[:, 0], boundary, nx)
    if mode == "linear":
        return Table(np.interp(x, data[:, 0], data[:, 1]))
    if mode == "cubic":
        from scipy.interpolate import CubicSpline

        return Table(CubicSpline(data[:, 0], data[:, 1])(x))
    raise RuntimeError("Unknown interpolation mode!")


def sampleTable2D(data, nx: int = 1024, ny: int = 1024, *, boundaries=None,
                  mode: Literal["linear", "cubic"] = "linear") -> Table:
    """Resamples scattered (x, y, f(x,y)) rows onto a regular nx x ny grid."""
    from scipy.interpolate import CloughTocher2DInterpolator, LinearNDInterpolator

    data = np.asarray(data)
    if boundaries is None:
        bx = by = None
    elif type(boundaries) is tuple and len(boundaries) == 2:
        bx, by = boundaries
    else:
        raise RuntimeError("Cant parse given boundaries!")
    x, y = np.meshgrid(_grid(data[:, 0], bx, nx), _grid(data[:, 1], by, ny))
    if mode == "linear":
        model = LinearNDInterpolator
    elif mode == "cubic":
        model = CloughTocher2DInterpolator
    else:
        raise RuntimeError("Unknown interpolation mode!")
    return Table(model(data[:, :2], data[:, 2])(x, y))


def evalTable(f, *ai) -> Table:
    """
    Samples ``f`` on a regular grid; each axis is ``n`` (on [0,1]) or ``(min, max, n)``.
    (The reference's version raises on ``axes.push``, src/theia/lookup.py:273; this one
    returns the table of function values.)
    """

    def axis(spec):
        if type(spec) is int:
            return np.linspace(0.0, 1.0, spec)
        if type(spec) is tuple and len(spec) == 3:
            return np.linspace(*spec)
        raise RuntimeError(f"Cannot parse dimension: {spec}")

    grid = np.meshgrid(*[axis(a) for a in ai], indexing="ij")
    return Table(f(*grid))
