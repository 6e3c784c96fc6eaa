"""
Structure-of-arrays queues shared between host and device: stand-in for the part of
``hephaistos.queue`` theia's responses use (IOQueue, QueueView).

Memory layout (sh/response.queue.glsl:6-33, sh/response.record.glsl:6-9): ``uint32 count``
followed by one array of ``capacity`` elements per scalar component, in field order
(``posX[C], posY[C], posZ[C], dirX[C], ...``).
"""

from __future__ import annotations

import ctypes
from ctypes import Structure

import numpy as np
import torch

from theia_b200 import _device
from theia_b200.pipeline import Command

__all__ = ["IOQueue", "QueueView", "queueSize"]

_NP = {ctypes.c_float: np.float32, ctypes.c_int32: np.int32, ctypes.c_uint32: np.uint32}


def _components(item: type[Structure]) -> list[tuple[str, type, int]]:
    out = []
    for name, t in item._fields_:
        if issubclass(t, ctypes.Array):
            out.append((name, _NP[t._type_], t._length_))
        else:
            out.append((name, _NP[t], 1))
    return out


def queueSize(item: type[Structure], capacity: int, *, skipCounter: bool = False) -> int:
    """Bytes needed for a queue of the given item type and capacity."""
    return (0 if skipCounter else 4) + 4 * capacity * sum(n for _, _, n in _components(item))


class QueueView:
    """
    View on a queue in host memory: ``view.count``, ``view["field"]`` (shape ``(n,)`` or
    ``(n, k)``), sliceable and indexable along the item axis like the reference's QueueView.
    """

    def __init__(self, data: np.ndarray | None, item: type[Structure], capacity: int, *, skipCounter: bool = False,
                 _fields=None, _count=None):
        self._item = item
        self._capacity = capacity
        self._header = None
        if _fields is not None:
            self._fields = _fields
            self._fixedCount = _count
            return
        self._fixedCount = None
        self._fields = {}
        if skipCounter:  # queues that are always full (host sources): no count word
            self._fixedCount = capacity
            off = 0
        else:
            self._header = data[:4].view(np.uint32)
            off = 4
        for name, dt, n in _components(item):
            arr = data[off : off + 4 * capacity * n].view(dt)
            self._fields[name] = arr if n == 1 else arr.reshape(n, capacity).T
            off += 4 * capacity * n

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def count(self) -> int:
        if self._header is None:
            return self._fixedCount
        return int(self._header[0])

    @count.setter
    def count(self, value: int) -> None:
        if self._header is None:
            raise RuntimeError("cannot set the count of a queue slice")
        self._header[0] = value

    @property
    def fields(self) -> list[str]:
        return list(self._fields)

    def __len__(self) -> int:
        return self._capacity

    def __getitem__(self, key):
        if isinstance(key, str):
            return self._fields[key]
        sub = {n: a[key] for n, a in self._fields.items()}
        n = len(next(iter(sub.values())))
        return QueueView(None, self._item, n, _fields=sub, _count=n)

    def __setitem__(self, key: str, value) -> None:
        self._fields[key][...] = value


class IOQueue:
    """
    Queue living on the device with two host staging copies, one per pipeline configuration
    (hephaistos.queue.IOQueue). ``mode="retrieve"``: device -> host after the run and the
    device count is cleared; ``mode="update"``: host -> device before the run.
    """

    def __init__(self, item: type[Structure], capacity: int = 0, *, mode: str = "retrieve",
                 skipCounter: bool = False) -> None:
        if mode not in ("retrieve", "update"):
            raise ValueError("mode must be 'retrieve' or 'update'")
        self._item = item
        self._mode = mode
        self._skipCounter = skipCounter
        self._capacity = 0
        self._tensor = None
        self._buffers = None
        self._device = _device.currentDevice()
        if capacity > 0:
            self.initialize(capacity)

    @property
    def initialized(self) -> bool:
        return self._tensor is not None

    @property
    def capacity(self) -> int:
        return self._capacity

    @property
    def item(self) -> type[Structure]:
        return self._item

    @property
    def tensor(self) -> torch.Tensor:
        """Device buffer holding the queue"""
        return self._tensor

    def initialize(self, capacity: int) -> None:
        if self.initialized:
            raise RuntimeError("queue is already initialized")
        self._capacity = capacity
        n = queueSize(self._item, capacity, skipCounter=self._skipCounter)
        self._nbytes = n
        self._tensor = _device.alloc(n, self._device)
        self._buffers = [_device.allocPinned(n) for _ in range(2)]
        if self._mode == "update" and not self._skipCounter:
            for i in range(2):
                self.view(i).count = capacity

    def buffer(self, i: int) -> torch.Tensor:
        return self._buffers[i]

    def view(self, i: int) -> QueueView:
        """View of the i-th host copy"""
        return QueueView(self._buffers[i].numpy(), self._item, self._capacity, skipCounter=self._skipCounter)

    def run(self, i: int) -> list[Command]:
        if self._mode == "retrieve":
            def cmd() -> None:
                self._buffers[i].copy_(self._tensor, non_blocking=True)
                if not self._skipCounter:
                    self._tensor[:4].zero_()  # reset count for the next batch
        else:
            def cmd() -> None:
                self._tensor.copy_(self._buffers[i], non_blocking=True)
        return [cmd]
