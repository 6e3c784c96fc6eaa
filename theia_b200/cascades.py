"""
Particle -> light source parameters (API of the reference's src/theia/cascades.py): the host side
step in front of the tracers that turns a particle of an event generator into the parameters of a
``ParticleCascadeLightSource`` or ``MuonTrackLightSource``. Nothing here runs on the device.

The numbers are the fits of L. Raedel's Geant4 studies the reference uses (arXiv:1206.5530,
arXiv:1210.5140; src/theia/cascades.py:203-344), held as one table instead of one object per
particle. Two peculiarities of the reference are kept because its outputs are what has to be
reproduced (tests/golden/reference_host.npz holds them):
  * ``createCascadeParameters`` computes both angular parameters from ``a_angluar_shift``
    (src/theia/cascades.py:187-188); the slopes and ``b_angular_shift`` of the fits never enter;
  * ``createParamsFromParticle`` hands ``x0`` / ``density`` on under names the converters do not
    read (src/theia/cascades.py:507-512 vs :431-437), so cascades always use sea water.
"""

from __future__ import annotations

from dataclasses import dataclass
from enum import IntEnum
from typing import Final

import numpy as np
from scipy.stats import norm

import theia_b200.units as u
from theia_b200.light import LightSource, MuonTrackLightSource, ParticleCascadeLightSource

__all__ = [
    "CascadeLightYield", "CascadeParameters", "CascadePrimaryParticle", "EMinus", "EPlus", "Gamma", "K0_Long",
    "Neutron", "Particle", "ParticleType", "PiMinus", "PiPlus", "PMinus", "PPlus", "X0_ice", "X0_water",
    "createCascadeParameters", "createParamsFromParticle", "getCascadeParamsFromParticleType", "rho_ice",
    "rho_water",
]

X0_ice: Final[float] = 39.75 * u.cm
"""Radiation length in ice"""
X0_water: Final[float] = 36.08 * u.cm
"""Radiation length in water"""
rho_ice: Final[float] = 0.91
"""Density of ice [g/cm3]"""
rho_water: Final[float] = 1.039
"""Density of deep sea water [g/cm3]"""


class ParticleType(IntEnum):
    """PDG Monte Carlo particle numbers (src/theia/cascades.py:70-88)"""

    UNKNOWN = 0
    GAMMA = 22
    E_PLUS = -11
    E_MINUS = 11
    MU_PLUS = -13
    MU_MINUS = 13
    TAU_PLUS = -15
    TAU_MINUS = 15
    PI_0 = 111
    PI_PLUS = 211
    PI_MINUS = -211
    K0_LONG = 130
    NEUTRON = 2112
    P_PLUS = 2212
    P_MINUS = -2212


@dataclass
class Particle:
    """A particle as an event generator hands it over (src/theia/cascades.py:91-109)"""

    particleType: ParticleType
    position: tuple[float, float, float]
    direction: tuple[float, float, float]
    energy: float
    startTime: float = 0.0
    length: float = float("NaN")  # track length where applicable
    speed: float = 1.0 * u.c


@dataclass
class CascadeLightYield:
    """Length of the Cherenkov track with the same light yield, and its fluctuation"""

    effectiveLength: float
    effectiveLengthStd: float = 0.0


@dataclass
class CascadeParameters:
    """Parameters of the cascade light source (longitudinal gamma profile, yield, angular profile)"""

    a_long: float
    b_long: float
    effectiveLength: float
    effectiveLengthStd: float = 0.0
    a_angular: float = 0.5375
    b_angular: float = 3.302


@dataclass(frozen=True)
class CascadePrimaryParticle:
    """Fit results for one kind of primary particle; every parameter is p = alpha + beta log10(E) or alpha E^beta"""

    alpha_long: float
    beta_long: float
    b_long: float
    alpha_length: float = 5.321
    beta_length: float = 1.0
    alpha_length_std: float = 5.727e-2
    beta_length_std: float = 0.5
    a_angluar_shift: float = 0.5375  # (sic: the field name of the reference)
    a_angular_slope: float = 0.0
    b_angular_shift: float = 3.302
    b_angular_slope: float = 0.0


# One row per primary: alpha_long, beta_long, b_long, alpha_length, beta_length, alpha_length_std,
# beta_length_std, a_angular shift / slope, b_angular shift / slope (src/theia/cascades.py:203-344)
_FITS = {
    "EMinus": (2.01849, 1.45469, 0.63207, 5.3207078881, 1.00000211, 0.0578170887, 0.5,
               0.53734995, 0.0, 3.30382993, 0.0),
    "EPlus": (2.00035, 1.45501, 0.63008, 5.3211320598, 0.99999254, 0.0573419669, 0.5,
              0.5367158, 0.0, 3.30484209, 0.0),
    "Gamma": (2.83923, 1.45501, 0.64526, 5.3208540905, 0.99999877, 0.0566586567, 0.5,
              0.53841841, 0.0, 3.29619817, 0.0),
    "PiPlus": (1.58357292, 0.96447937, 0.33833116, 3.3355182722, 1.03662217, 1.1920455395, 0.80772057,
               1.0299732199972658, -0.08806219920032332, 3.102713004779744, -0.12229465620485062),
    "PiMinus": (1.69176636, 0.93953506, 0.34108075, 3.3584489578, 1.03584394, 1.2250188073, 0.80322520,
                1.0412256610000645, -0.09187703681909758, 3.086039699134421, -0.11874011144663844),
    "K0_Long": (1.95948974, 0.80440041, 0.34535151, 3.2600450524, 1.03931457, 1.2141970572, 0.80779629,
                1.0591474180300977, -0.09635256670474648, 3.2258115113151793, -0.15816716921465757),
    "PPlus": (1.92249171, 0.77601150, 0.34969748, 2.8737183922, 1.05172118, 0.8804581378, 0.82445572,
              1.1574216500437113, -0.11090280215147694, 3.5079727644060794, -0.22892116764330248),
    "PMinus": (1.92249171, 0.77601150, 0.34969748, 3.0333074914, 1.04322206, 1.1323088104, 0.77134060,
               1.1574216500437113, -0.11090280215147694, 3.5079727644060794, -0.22892116764330248),
    "Neutron": (1.57739060, 0.93556570, 0.35269455, 2.7843854660, 1.05582906, 0.9322787137, 0.81776503,
                1.1292267334081203, -0.10876633838986713, 3.4157386880981093, -0.20638832466150736),
}
EMinus, EPlus, Gamma, PiPlus, PiMinus, K0_Long, PPlus, PMinus, Neutron = (
    CascadePrimaryParticle(*_FITS[k]) for k in ("EMinus", "EPlus", "Gamma", "PiPlus", "PiMinus", "K0_Long", "PPlus",
                                                 "PMinus", "Neutron"))

_BY_TYPE = {
    ParticleType.GAMMA: Gamma, ParticleType.E_MINUS: EMinus, ParticleType.E_PLUS: EPlus,
    ParticleType.PI_0: Gamma,  # decays at once into two photons
    ParticleType.PI_PLUS: PiPlus, ParticleType.PI_MINUS: PiMinus, ParticleType.K0_LONG: K0_Long,
    ParticleType.P_PLUS: PPlus, ParticleType.P_MINUS: PMinus, ParticleType.NEUTRON: Neutron,
}
_TRACKS = {ParticleType.MU_PLUS, ParticleType.MU_MINUS, ParticleType.TAU_PLUS, ParticleType.TAU_MINUS}


def getCascadeParamsFromParticleType(t: ParticleType) -> CascadePrimaryParticle | None:
    """Fits of the cascade a particle of type ``t`` starts, or None if it does not start one"""
    return _BY_TYPE.get(t)


def createCascadeParameters(p: CascadePrimaryParticle, E: float, X0: float = X0_water,
                            density: float = rho_water) -> CascadeParameters:
    """Cascade parameters of primary ``p`` with energy ``E`` [GeV] (src/theia/cascades.py:165-200)"""
    logE = max(0.0, np.log10(E))
    a_long = p.alpha_long + p.beta_long * logE
    b_long = X0 / p.b_long
    # as the reference computes them (see the module docstring)
    a_angular = p.a_angluar_shift * logE + p.a_angluar_shift
    b_angular = p.b_angular_slope * logE + p.a_angluar_shift
    rhoScale = 0.91 / density  # eq. (9) of arXiv:1210.5140
    effectiveLength = p.alpha_length * rhoScale * (E**p.beta_length)
    effectiveLengthStd = p.alpha_length_std * rhoScale * (E**p.beta_length_std)
    return CascadeParameters(a_long, b_long, effectiveLength, effectiveLengthStd, a_angular, b_angular)


def _prefix(name: str) -> str:
    return name + "__" if name else name


def _trackParams(particle: Particle, name: str, uRand: float | None):
    """Muon like particles -> MuonTrackLightSource (src/theia/cascades.py:380-428)"""
    if particle.particleType not in _TRACKS:
        return None
    if not particle.length > 0.0:  # also catches NaN
        raise ValueError("particle is muon like, but no track length was specified!")
    pos, d = np.asarray(particle.position, dtype=np.float64), np.asarray(particle.direction, dtype=np.float64)
    end = pos + d * (particle.length / np.sqrt(np.square(d).sum()))
    scale = 1.1880 + 0.0206 * np.log(particle.energy)
    length = particle.length * scale
    std = np.sqrt(particle.length * 0.1 * u.m) * scale  # eq. 7 of arXiv:1206.5530
    if uRand is not None:
        # at least the light of the bare muon
        length = max(length + norm.ppf(uRand).item() * std, particle.length)
        std = 0.0
    n = _prefix(name)
    params = {f"{n}startPosition": particle.position, f"{n}startTime": particle.startTime,
              f"{n}endPosition": tuple(float(x) for x in end),
              f"{n}endTime": particle.startTime + particle.length / particle.speed, f"{n}muonEnergy": particle.energy}
    return MuonTrackLightSource, params, CascadeLightYield(length, std)


def _cascadeParams(particle: Particle, name: str, uRand: float | None):
    """Everything that showers -> ParticleCascadeLightSource (src/theia/cascades.py:431-478)"""
    primary = getCascadeParamsFromParticleType(particle.particleType)
    if primary is None:
        return None
    c = createCascadeParameters(primary, particle.energy, X0_water, rho_water)
    length, std = c.effectiveLength, c.effectiveLengthStd
    if uRand is not None:
        length = max(0.0, length + norm.ppf(uRand).item() * std)
        std = 0.0
    d = np.asarray(particle.direction, dtype=np.float64)
    d = d / np.sqrt(np.square(d).sum())
    n = _prefix(name)
    params = {f"{n}startPosition": particle.position, f"{n}startTime": particle.startTime,
              f"{n}direction": tuple(float(x) for x in d), f"{n}effectiveLength": length,
              f"{n}a_angular": c.a_angular, f"{n}b_angular": c.b_angular, f"{n}a_long": c.a_long, f"{n}b_long": c.b_long}
    return ParticleCascadeLightSource, params, CascadeLightYield(length, std / length)


def createParamsFromParticle(particle: Particle, *, x0: float = X0_water, density: float = rho_water,
                             lightSourceName: str = "lightSource", uRand: float | None = None
                             ) -> tuple[type[LightSource], dict, CascadeLightYield]:
    """
    Light source class, its stage parameters (``<lightSourceName>__<param>``, ready for a pipeline
    task) and the light yield of ``particle``; ``uRand`` in (0, 1) samples the fluctuation of the
    yield instead of reporting it (src/theia/cascades.py:481-530).
    """
    for convert in (_trackParams, _cascadeParams):
        res = convert(particle, lightSourceName, uRand)
        if res is not None:
            return res
    raise ValueError(f"Could not create params from particle '{particle}'!")
