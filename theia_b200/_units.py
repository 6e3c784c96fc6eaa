"""
Host side of the component / shader unit seam (``theia_unit_eval``, include/theia_b200.h): evaluates
one function of the photon path for ``n`` inputs on the device. It is what the sampler stages
(``light.LightSampler``, ``testing.TargetSampler``, ``testing.TargetGuideSampler``) run, and what the
ports of the reference's shader unit tests call (``tests/test_reference_units.py``).
"""

from __future__ import annotations

from ctypes import byref, c_uint32, sizeof

import numpy as np
import torch

from theia_b200 import _abi, _device, _lib

__all__ = ["unitArgs", "unitWidths", "evalUnit"]

#: row widths (floats in, floats out) per unit; mirrors theia_unit_widths
WIDTHS = {
    _abi.UNIT_SCATTER_DIR: (5, 3), _abi.UNIT_REFLECTANCE: (7, 3), _abi.UNIT_VOLUME_SCATTER: (5, 4),
    _abi.UNIT_SCATTER_PROB: (8, 1), _abi.UNIT_ROTATE_PHI: (5, 4), _abi.UNIT_ALIGN_POLREF: (0, 16),
    _abi.UNIT_ROTATE_DIR: (12, 7), _abi.UNIT_PHASE_MATRIX: (5, 4), _abi.UNIT_SURFACE: (0, 33),
    _abi.UNIT_SIGN_BIT: (1, 1), _abi.UNIT_GAMMA: (0, 1), _abi.UNIT_EMISSION_ANGLE: (1, 1),
    _abi.UNIT_TARGET_SAMPLE: (0, 29), _abi.UNIT_GUIDE_SAMPLE: (0, 13), _abi.UNIT_LIGHT_SAMPLE: (0, 0),
    _abi.UNIT_MEDIUM: (3, 12),
}


def unitWidths(unit: int, polarized: bool = False) -> tuple[int, int]:
    """Row widths as the library reports them (checked against WIDTHS by the ABI tests)"""
    wi, wo = c_uint32(), c_uint32()
    _lib.check(_lib.load().theia_unit_widths(unit, int(polarized), byref(wi), byref(wo)))
    return wi.value, wo.value


def unitArgs(unit: int, n: int, *, rng=None, i: int = 0, polarized: bool = False, first: int = 0, baseCount: int = 0,
             medium: int = 0, material: int = 0, params=(), light=None, wavelength=None, target=None, guide=None,
             keep: list | None = None) -> _abi.UnitArgs:
    """Launch description of one unit; components are pipeline stages whose configuration ``i`` is committed."""
    from theia_b200.trace import _component
    keep = keep if keep is not None else []
    A = _abi.UnitArgs()
    A.struct_size = sizeof(_abi.UnitArgs)
    A.unit, A.n, A.polarized = unit, n, int(polarized)
    if rng is not None:
        A.rng = rng._abiParams(i)
    A.first, A.base_count = first, baseCount
    A.medium, A.material = int(medium or 0), int(material or 0)
    for k, v in enumerate(params):
        A.params[k] = float(v)
    A.light = _component(light, i, keep)
    A.wavelength = _component(wavelength, i, keep)
    A.target = _component(target, i, keep)
    A.guide = _component(guide, i, keep)
    A._keep = keep  # parameter blocks must outlive the call
    return A


def evalUnit(unit: int, rows: np.ndarray | None, n: int | None = None, **kw) -> np.ndarray:
    """
    Evaluates ``unit`` on the CUDA device for the input rows (float32 ``[n, in_width]``, or None for units
    without input) and returns the output rows as float32 ``[n, out_width]``.
    """
    if not torch.cuda.is_available():
        raise RuntimeError("theia_unit_eval needs a CUDA device (no CPU fallback)")
    wi, wo = WIDTHS[unit]
    if rows is not None:
        rows = np.ascontiguousarray(rows, dtype=np.float32).reshape(-1, wi)
        n = len(rows)
    A = unitArgs(unit, int(n), **kw)
    tin = torch.from_numpy(rows).cuda() if rows is not None else None
    tout = torch.zeros((int(n), wo), dtype=torch.float32, device="cuda")
    A.in_ = tin.data_ptr() if tin is not None else None
    A.out = tout.data_ptr()
    A.stream = _device.currentStream("cuda")
    _lib.check(_lib.load().theia_unit_eval(byref(A)))
    return tout.cpu().numpy()
