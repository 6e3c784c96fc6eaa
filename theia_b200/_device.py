"""
Device selection and raw buffers. PyTorch is plumbing here: it owns device memory, pinned
host memory and streams; all compute goes through the C-ABI library (theia_b200._lib).

Buffers are torch uint8 tensors; "device addresses" handed to parameter blocks are
``tensor.data_ptr()`` values (the reference hands out raw Vulkan buffer addresses the same
way: src/theia/material.py:956-972, src/theia/scene.py:491-499).

The "cpu" device exists for host-side logic tests and for the parity harness, which builds
the same launch description over host memory and hands it to the CPU oracle. Tracers refuse
to run on it: there is no CPU fallback in the product path.
"""

from __future__ import annotations

import contextlib
import threading

import numpy as np
import torch

__all__ = ["alloc", "allocPinned", "currentDevice", "useDevice", "asNumpy", "currentStream"]

_local = threading.local()


def _default() -> str:
    return "cuda" if torch.cuda.is_available() else "cpu"


def currentDevice() -> str:
    """Device new objects are created on: "cuda" or "cpu"."""
    return getattr(_local, "device", None) or _default()


@contextlib.contextmanager
def useDevice(device: str):
    """Context manager selecting the device new objects are created on."""
    if device not in ("cuda", "cpu"):
        raise ValueError("device must be 'cuda' or 'cpu'")
    if device == "cuda" and not torch.cuda.is_available():
        raise RuntimeError("CUDA device requested but not available")
    old = getattr(_local, "device", None)
    _local.device = device
    try:
        yield
    finally:
        _local.device = old


def alloc(nbytes: int, device: str | None = None, *, zero: bool = True) -> torch.Tensor:
    """Allocates a raw byte buffer (16 byte granular) on the given device."""
    device = device or currentDevice()
    n = max(16, (int(nbytes) + 15) // 16 * 16)
    fn = torch.zeros if zero else torch.empty
    return fn(n, dtype=torch.uint8, device=device)


def allocPinned(nbytes: int) -> torch.Tensor:
    """Allocates a host staging buffer, pinned if CUDA is available."""
    n = max(16, (int(nbytes) + 15) // 16 * 16)
    pin = torch.cuda.is_available()
    return torch.zeros(n, dtype=torch.uint8, pin_memory=pin)


def asNumpy(t: torch.Tensor, dtype=np.uint8) -> np.ndarray:
    """Numpy view of a host tensor."""
    return t.numpy().view(dtype)


def currentStream(device: str) -> int:
    """cudaStream_t of torch's current stream (0 on the cpu device)."""
    if device == "cuda":
        return torch.cuda.current_stream().cuda_stream
    return 0
