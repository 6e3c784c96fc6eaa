"""
Generates tests/golden/reference_host.npz by importing the host-only numerics of the
reference itself (needs /root/reference: run in the build container, not on the GPU box).

    python tests/golden/generate_reference_vectors.py

The reference's modules import `hephaistos` (Vulkan runtime, absent) and `trimesh` at module
level; both are replaced by empty stand-ins, which is enough for the pure numpy / scipy code
used here: the medium models (src/theia/material.py), the test water model
(src/theia/testing.py:641-656), the material flag parser and scene.Transform. The arrays
written are exactly what the reference would upload to the GPU as look-up tables.
"""

import ctypes
import importlib.abc
import importlib.machinery
import os
import sys
import types

import numpy as np

REFERENCE = os.environ.get("THEIA_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_host.npz")


class _Stub(types.ModuleType):
    __path__ = []

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        if self.__name__ == "hephaistos.glsl":  # vec3, mat4x3, ... are used as ctypes field types
            return type(name, (ctypes.Structure,), {"_fields_": [("v", ctypes.c_float * 16)]})
        return type(name, (object,), {"__init__": lambda s, *a, **k: None,
                                      "__class_getitem__": classmethod(lambda c, i: c)})


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path, target=None):
        if fullname.split(".")[0] in ("hephaistos", "trimesh"):
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        return _Stub(spec.name)

    def exec_module(self, module):
        pass


sys.meta_path.insert(0, _StubFinder())
sys.path.insert(0, os.path.join(REFERENCE, "src"))

import theia.material as tm  # noqa: E402
import theia.scene as ts  # noqa: E402
import theia.testing as tt  # noqa: E402

TABLES = ["refractive_index", "group_velocity", "absorption_coef", "scattering_coef", "log_phase_function",
          "phase_sampling", "phase_m12", "phase_m22", "phase_m33", "phase_m34"]


class Homogenous(tm.DispersionFreeMedium, tm.HenyeyGreensteinPhaseFunction, tm.MediumModel):
    def __init__(self):
        tm.DispersionFreeMedium.__init__(self, n=1.33, ng=1.33, mu_a=0.05, mu_s=0.02)
        tm.HenyeyGreensteinPhaseFunction.__init__(self, 0.2)


class Energy(tm.DispersionFreeMedium, tm.HenyeyGreensteinPhaseFunction, tm.KokhanovskyOceanWaterPhaseMatrix,
             tm.MediumModel):
    def __init__(self):
        tm.DispersionFreeMedium.__init__(self, n=1.33, ng=1.33, mu_a=0.05, mu_s=0.01)
        tm.HenyeyGreensteinPhaseFunction.__init__(self, 0.9)
        tm.KokhanovskyOceanWaterPhaseMatrix.__init__(self, p90=0.66, theta0=0.25, alpha=4.0, xi=25.6)


class FF(tm.DispersionFreeMedium, tm.FournierForandPhaseFunction, tm.MediumModel):
    def __init__(self):
        tm.DispersionFreeMedium.__init__(self, n=1.35, ng=1.38, mu_a=0.03, mu_s=0.1)
        tm.FournierForandPhaseFunction.__init__(self, 1.175, 4.065)


out = {}
for name, model in {"water": tt.WaterTestModel(), "bk7": tm.BK7Model(), "homogenous": Homogenous(),
                    "energy": Energy(), "ff": FF()}.items():
    med = model.createMedium(name=name)
    out[f"{name}.range"] = np.array([med.lambda_min, med.lambda_max], np.float64)
    for t in TABLES:
        v = getattr(med, t)
        if v is not None:
            out[f"{name}.{t}"] = np.asarray(v, np.float64)

flagStrings = ["", "T", "R", "TR", "B", "D", "DB", "BD", "L", "V", "Tf", "Tb", "Rf", "Rb", "TfRb", "VD", "TRD"]
out["flags.strings"] = np.array(flagStrings)
out["flags.values"] = np.array([int(tm.parseMaterialFlags(s)) for s in flagStrings], np.int64)

T = ts.Transform
xf = [T.TRS(scale=0.8, translate=(3.0, 0.0, 0.0)),
      T.TRS(scale=(0.1, 5.0, 5.0), rotate=(0.0, 0.0, 1.0, 45.0), translate=(0.141, 0.0, 0.0)),
      T.TRS(scale=2.0, rotate=(1.0, 0.0, 0.0, 90.0)),
      T.TRS(scale=(1.0, 2.0, 3.0), rotate=(1.0, 2.0, -0.5, 33.0), translate=(-4.0, 5.5, 0.25))]
out["transform.trs"] = np.stack([np.asarray(t.numpy(), np.float64) for t in xf])

# theia.cascades: particle -> light source parameters (host side tables and formulas)
import theia.cascades as tc  # noqa: E402

CASC_TYPES = [22, -11, 11, 111, 211, -211, 130, 2112, 2212, -2212]
CASC_ENERGIES = [0.5, 1.0, 7.5, 100.0, 3.3e4]
rows = []
for t in CASC_TYPES:
    for E in CASC_ENERGIES:
        c = tc.createCascadeParameters(tc.getCascadeParamsFromParticleType(tc.ParticleType(t)), E)
        c2 = tc.createCascadeParameters(tc.getCascadeParamsFromParticleType(tc.ParticleType(t)), E, tc.X0_ice, tc.rho_ice)
        rows.append([c.a_long, c.b_long, c.effectiveLength, c.effectiveLengthStd, c.a_angular, c.b_angular,
                     c2.b_long, c2.effectiveLength])
out["cascades.parameters"] = np.array(rows, np.float64)
rows = []
for t, E, uRand in [(11, 50.0, None), (211, 800.0, 0.8), (22, 3.0, 0.1), (13, 120.0, None), (-15, 2e3, 0.35)]:
    p = tc.Particle(tc.ParticleType(t), (1.0, -2.0, 3.5), (0.3, -0.4, 1.2), E, startTime=12.5, length=75.0)
    cls, params, ly = tc.createParamsFromParticle(p, lightSourceName="src", uRand=uRand)
    flat = []
    for k in sorted(params):
        flat += list(np.atleast_1d(np.asarray(params[k], np.float64)))
    rows.append(([1.0] if cls is tc.MuonTrackLightSource else [0.0]) + flat + [ly.effectiveLength, ly.effectiveLengthStd] +
                [0.0] * 8)
out["cascades.from_particle"] = np.array([r[:16] for r in rows], np.float64)

np.savez_compressed(OUT, **out)
print(f"wrote {OUT}: {len(out)} arrays, {os.path.getsize(OUT)} bytes")
