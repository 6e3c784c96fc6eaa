"""
Loader for the C-ABI library ``libtheia_b200.so`` (include/theia_b200.h). The library is
built in-tree by ``__graft_entry__.build()`` / ``theia_b200/csrc/build.py``. If it is missing
every entry point raises: there is no fallback path.
"""

from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int32, c_uint32, c_uint64, c_void_p

from theia_b200 import _abi

# THEIA_B200_LIB selects another build of the same library (A/B measurements of kernel variants)
_LIB_PATH = os.environ.get("THEIA_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)),
                                                             "libtheia_b200.so")
_lib = None

#: every symbol include/theia_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "theia_last_error": (c_char_p, []),
    "theia_abi_version": (c_int, []),
    "theia_kernel_launches": (c_uint64, []),
    "theia_set_option": (c_int, [c_char_p, c_int]),
    "theia_get_trace_choice": (c_int, [c_void_p, c_uint32, c_int, POINTER(c_int), POINTER(c_int)]),
    "theia_profile_enable": (c_int, [c_int]),
    "theia_profile_read": (c_int, [POINTER(ctypes.c_double), POINTER(c_uint64), c_uint32]),
    "theia_profile_class_name": (c_char_p, [c_uint32]),
    "theia_trace_stats_enable": (c_int, [c_void_p]),
    "theia_l2_probe": (c_int, [c_void_p, c_uint64, c_uint32, c_void_p, c_void_p]),
    "theia_device_info": (c_int, [POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    "theia_philox_fill": (c_int, [POINTER(_abi.PhiloxParams), c_uint32, c_uint32, c_uint32, c_uint32, c_uint32,
                                  c_void_p, c_void_p]),
    "theia_philox_raw": (c_int, [c_void_p, c_void_p, c_uint32, c_void_p, c_void_p]),
    "theia_blas_sizes": (c_int, [c_uint32, POINTER(_abi.ASSizes)]),
    "theia_blas_build": (c_int, [c_void_p, c_uint32, c_void_p, c_uint32, c_void_p, c_uint64, c_void_p, c_uint64,
                                 c_void_p]),
    "theia_tlas_sizes": (c_int, [c_uint32, POINTER(_abi.ASSizes)]),
    "theia_tlas_build": (c_int, [POINTER(_abi.InstanceDesc), c_uint32, c_void_p, c_uint64, c_void_p, c_uint64,
                                 c_void_p]),
    "theia_tlas_query": (c_int, [c_void_p, c_void_p, c_uint32, c_void_p, c_void_p, c_void_p, c_void_p]),
    "theia_tlas_query_static": (c_int, [c_void_p, c_void_p, c_uint32, c_void_p, c_void_p, c_void_p]),
    "theia_trace_grid_size": (c_int, [POINTER(_abi.TraceLaunch), POINTER(c_uint32)]),
    "theia_trace_workspace_size": (c_int, [POINTER(_abi.TraceLaunch), POINTER(c_uint64)]),
    "theia_trace_forward": (c_int, [POINTER(_abi.TraceLaunch)]),
    "theia_hist_reduce": (c_int, [c_void_p, c_uint32, c_uint32, c_float, c_void_p, c_int, c_void_p]),
    "theia_hit_replay": (c_int, [c_void_p, c_uint32, c_uint32, POINTER(_abi.Response), c_uint32, c_void_p]),
    "theia_hit_replay_rng": (c_int, [c_void_p, c_uint32, c_uint32, POINTER(_abi.Response), c_void_p, c_uint32, c_void_p]),
    "theia_math_eval": (c_int, [c_uint32, c_void_p, c_void_p, c_uint32, c_void_p, c_void_p]),
    "theia_unit_widths": (c_int, [c_uint32, c_uint32, POINTER(c_uint32), POINTER(c_uint32)]),
    "theia_unit_eval": (c_int, [POINTER(_abi.UnitArgs)]),
}


def libraryPath() -> str:
    return _LIB_PATH


def load() -> ctypes.CDLL:
    """Loads the library (once). Raises RuntimeError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise RuntimeError(
            f"{_LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
            "theia_b200 has no CPU fallback."
        )
    lib = ctypes.CDLL(_LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.theia_abi_version() != _abi.ABI_VERSION:
        raise RuntimeError("libtheia_b200.so has a different ABI version than the Python package")
    _lib = lib
    return lib


def check(code: int) -> None:
    """Raises RuntimeError with the library's message if ``code`` signals an error."""
    if code != 0:
        msg = load().theia_last_error()
        raise RuntimeError(f"theia_b200 error {code}: {msg.decode() if msg else 'unknown'}")


def smCount() -> int:
    sm, major, minor = c_int(), c_int(), c_int()
    check(load().theia_device_info(ctypes.byref(sm), ctypes.byref(major), ctypes.byref(minor)))
    return sm.value
