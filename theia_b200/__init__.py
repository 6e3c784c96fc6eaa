"""
theia_b200 - B200-native implementation of the forward photon tracing path of
tkerscher/theia behind theia's own Python component API.

Submodules mirror the reference's module names (src/theia/__init__.py:21-36):
cascades, light, lookup, material, random, response, scene, target, task, testing, trace, units,
plus distributed (multi-GPU sharding).
All device work goes through the C-ABI library built from theia_b200/csrc
(include/theia_b200.h); there is no CPU fallback.
"""

import importlib

__all__ = [
    "cascades", "distributed", "light", "lookup", "material", "pipeline", "random", "response", "scene",
    "target", "task", "testing", "trace", "units",
]


def __dir__():
    return __all__


def __getattr__(attr):
    if attr in __all__:
        return importlib.import_module(f"theia_b200.{attr}")
    raise AttributeError(f"module 'theia_b200' has no attribute {attr!r}")
