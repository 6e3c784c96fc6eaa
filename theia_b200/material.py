"""
Media, materials and their device representation (API of the reference's
src/theia/material.py).

Device layouts (scalar, little endian):
 * Medium   (88 B, 8 aligned): float lambda_min, lambda_max; uint64 table addresses n, vg,
   mu_a, mu_s, log_phase, phase_sampling, phase_m12, phase_m22, phase_m33, phase_m34
   (0 = absent)                               - src/theia/material.py:122-140
 * Material (24 B, 8 aligned): uint64 inside, outside; uint32 flagsInwards, flagsOutwards
                                              - src/theia/material.py:587-598
 * Table: see theia_b200.lookup
``MaterialStore`` packs everything into ONE buffer and hands out raw addresses.
"""

from __future__ import annotations

import importlib.resources
import json
import re
import warnings
from collections.abc import Iterable
from ctypes import Structure, c_float, c_uint32, c_uint64, sizeof
from enum import IntFlag
from io import TextIOBase
from pathlib import Path
from types import MappingProxyType
from typing import Final
from zipfile import ZipFile, is_zipfile

import numpy as np
import torch
from numpy.typing import ArrayLike, NDArray

import theia_b200.units as u
from theia_b200 import _device
from theia_b200.lookup import Table, getTableSize

__all__ = [
    "loadMaterials", "parseMaterialFlags", "saveMaterials", "speed_of_light", "BK7Model",
    "DispersionFreeMedium", "FournierForandPhaseFunction", "HenyeyGreensteinPhaseFunction",
    "KokhanovskyOceanWaterPhaseMatrix", "Material", "MaterialFlags", "MaterialStore", "Medium",
    "MediumModel", "SellmeierEquation", "WaterBaseModel",
]

speed_of_light: Final[float] = 1.0 * u.c
"""speed of light in internal units (m/ns)"""

# (python property, table key suffix used by MaterialStore, GLSL field)
_TABLES = [
    ("refractive_index", "n", "n"),
    ("group_velocity", "vg", "vg"),
    ("absorption_coef", "mua", "mu_a"),
    ("scattering_coef", "mus", "mu_s"),
    ("log_phase_function", "lpf", "log_phase"),
    ("phase_sampling", "ps", "phase_sampling"),
    ("phase_m12", "m12", "phase_m12"),
    ("phase_m22", "m22", "phase_m22"),
    ("phase_m33", "m33", "phase_m33"),
    ("phase_m34", "m34", "phase_m34"),
]


class Medium:
    """
    Named medium: up to ten tables sampled at equidistant points. n, vg, mu_a, mu_s are
    functions of the vacuum wavelength on [lambda_min, lambda_max]; log_phase_function and
    phase_m12..m34 are functions of cos(theta) on [-1, 1]; phase_sampling is the inverse
    CDF of the phase function on [0, 1]. ``None`` means: n = 1, vg = c, mu = 0, isotropic
    scattering, no polarisation coupling. (src/theia/material.py:61-438)
    """

    class GLSL(Structure):
        _fields_ = [("lambda_min", c_float), ("lambda_max", c_float)] + [
            (glsl, c_uint64) for _, _, glsl in _TABLES
        ]

    ALIGNMENT: Final[int] = 8
    _PROPS = [p for p, _, _ in _TABLES]

    def __init__(self, name: str, lambda_min: float, lambda_max: float, *, refractive_index=None,
                 group_velocity=None, absorption_coef=None, scattering_coef=None, log_phase_function=None,
                 phase_sampling=None, phase_m12=None, phase_m22=None, phase_m33=None, phase_m34=None) -> None:
        self.name = name
        self.lambda_min = lambda_min
        self.lambda_max = lambda_max
        self.refractive_index = refractive_index
        self.group_velocity = group_velocity
        self.absorption_coef = absorption_coef
        self.scattering_coef = scattering_coef
        self.log_phase_function = log_phase_function
        self.phase_sampling = phase_sampling
        self.phase_m12 = phase_m12
        self.phase_m22 = phase_m22
        self.phase_m33 = phase_m33
        self.phase_m34 = phase_m34

    def save(self, file) -> None:
        """Saves the medium in numpy's .npz format (lambda_range + the non-empty tables)."""
        if isinstance(file, TextIOBase):
            raise ValueError("file must be opened in binary mode!")
        arrays = {p: getattr(self, p) for p in self._PROPS if getattr(self, p) is not None}
        arrays["lambda_range"] = np.array([self.lambda_min, self.lambda_max])
        np.savez(file, **arrays)

    @staticmethod
    def load(file, *, name: str = "unnamed") -> "Medium":
        if isinstance(file, TextIOBase):
            raise ValueError("file must be opened in binary mode!")
        data = np.load(file)
        lam = data.get("lambda_range")
        if lam is None or lam.shape != (2,):
            raise ValueError("File does not contain valid lambda range!")
        medium = Medium(name, lam[0], lam[1])
        for prop in Medium._PROPS:
            setattr(medium, prop, data.get(prop))
        return medium


class MaterialFlags(IntFlag):
    """Bit flags for what a ray does at a material (src/theia/material.py:441-510)."""

    BLACK_BODY = 0x01
    DETECTOR = 0x02
    LIGHT_SOURCE = 0x04
    NO_REFLECT_FWD = 0x08
    NO_REFLECT_BWD = 0x10
    NO_REFLECT = 0x18
    NO_TRANSMIT_FWD = 0x20
    NO_TRANSMIT_BWD = 0x40
    NO_TRANSMIT = 0x60
    VOLUME_BORDER = 0x80


_flagTokens = {
    "B": MaterialFlags.BLACK_BODY, "D": MaterialFlags.DETECTOR, "L": MaterialFlags.LIGHT_SOURCE,
    "R": MaterialFlags.NO_REFLECT, "Rbf": MaterialFlags.NO_REFLECT, "Rfb": MaterialFlags.NO_REFLECT,
    "Rb": MaterialFlags.NO_REFLECT_BWD, "Rf": MaterialFlags.NO_REFLECT_FWD,
    "T": MaterialFlags.NO_TRANSMIT, "Tbf": MaterialFlags.NO_TRANSMIT, "Tfb": MaterialFlags.NO_TRANSMIT,
    "Tb": MaterialFlags.NO_TRANSMIT_BWD, "Tf": MaterialFlags.NO_TRANSMIT_FWD,
    "V": MaterialFlags.VOLUME_BORDER,
}


def parseMaterialFlags(flags: str) -> MaterialFlags:
    """
    Parses a flag string: capital letters B, D, L, R, T, V, with R/T optionally narrowed by
    f/b. Starts from NO_REFLECT | NO_TRANSMIT and XORs each token, so "R"/"T" *enable*
    reflection/transmission (src/theia/material.py:531-557).
    """
    result = MaterialFlags.NO_REFLECT | MaterialFlags.NO_TRANSMIT
    for token in re.findall(r"[A-Z][a-z]*", flags):
        if token not in _flagTokens:
            raise ValueError(f"Unknown material flag '{token}'")
        result ^= _flagTokens[token]
    return result


class Material:
    """
    Material of a geometry: the media on the inside and outside (normals point outwards) and
    the flags for rays travelling inwards / outwards (src/theia/material.py:560-693).
    """

    class GLSL(Structure):
        _fields_ = [("inside", c_uint64), ("outside", c_uint64),
                    ("flagsInwards", c_uint32), ("flagsOutwards", c_uint32)]

    ALIGNMENT: Final[int] = 8

    def __init__(self, name: str, inside: Medium | str | None, outside: Medium | str | None, *,
                 flags=MaterialFlags(0)) -> None:
        self.name = name
        self.inside = inside
        self.outside = outside
        if type(flags) is tuple:
            self.flagsInward, self.flagsOutward = flags
        else:
            self.flagsInward = self.flagsOutward = flags

    @property
    def flagsInward(self) -> MaterialFlags:
        return self._flagsInward

    @flagsInward.setter
    def flagsInward(self, value) -> None:
        self._flagsInward = value if type(value) is MaterialFlags else parseMaterialFlags(value)

    @flagsInward.deleter
    def flagsInward(self) -> None:
        self._flagsInward = MaterialFlags(0)

    @property
    def flagsOutward(self) -> MaterialFlags:
        return self._flagsOutward

    @flagsOutward.setter
    def flagsOutward(self, value) -> None:
        self._flagsOutward = value if type(value) is MaterialFlags else parseMaterialFlags(value)

    @flagsOutward.deleter
    def flagsOutward(self) -> None:
        self._flagsOutward = MaterialFlags(0)


def _mediumName(m) -> str | None:
    if m is None:
        return None
    return m.name if isinstance(m, Medium) else str(m)


class MaterialStore:
    """
    Packs media, their tables and materials into one buffer and exposes their addresses by
    name (src/theia/material.py:884-1117). With ``freeze=False`` existing entries can be
    updated in the host copy and re-uploaded with ``flush()``.
    """

    def __init__(self, material: Iterable[Material], *, media: Iterable[Medium] = [], freeze: bool = True) -> None:
        material, media = list(material), list(media)
        self._frozen = False
        self._device = _device.currentDevice()
        size = 0

        def alloc(n: int, alignment: int) -> int:
            nonlocal size
            size = -(-size // alignment) * alignment
            ptr = size
            size += n
            return ptr

        self._tableOff: dict[str, tuple[int, int]] = {}
        self._mediaOff: dict[str, int] = {}
        self._matOff: dict[str, int] = {}

        def allocMedium(medium) -> None:
            name = _mediumName(medium)
            if name is None or name in self._mediaOff:
                return
            self._mediaOff[name] = alloc(sizeof(Medium.GLSL), Medium.ALIGNMENT)
            if isinstance(medium, Medium):
                for prop, key, _ in _TABLES:
                    data = getattr(medium, prop)
                    if data is not None:
                        n = getTableSize(data)
                        self._tableOff[f"{name}_{key}"] = (alloc(n, Table.ALIGNMENT), n)

        for m in media:
            allocMedium(m)
        for mat in material:
            if mat.name not in self._matOff:
                self._matOff[mat.name] = alloc(sizeof(Material.GLSL), Material.ALIGNMENT)
                allocMedium(mat.inside)
                allocMedium(mat.outside)

        self._size = size
        self._host = np.zeros(max(size, 16), dtype=np.uint8)
        self._tensor = _device.alloc(size, self._device)
        base = self._tensor.data_ptr()
        self._tableAdr = {k: base + off for k, (off, _) in self._tableOff.items()}
        self._mediaAdr = {k: base + off for k, off in self._mediaOff.items()}
        self._matAdr = {k: base + off for k, off in self._matOff.items()}
        self._media = MappingProxyType(self._mediaAdr)
        self._mat = MappingProxyType(self._matAdr)

        done: set[str] = set()

        def fill(medium) -> None:
            if isinstance(medium, Medium) and medium.name not in done:
                self.updateMedium(medium)
                done.add(medium.name)

        for m in media:
            fill(m)
        for mat in material:
            self.updateMaterial(mat, updateMedia=False)
            fill(mat.inside)
            fill(mat.outside)
        self.flush()
        self._frozen = freeze

    @property
    def frozen(self) -> bool:
        return self._frozen

    @property
    def material(self) -> MappingProxyType:
        """name -> device address of the material"""
        return self._mat

    @property
    def media(self) -> MappingProxyType:
        """name -> device address of the medium"""
        return self._media

    @property
    def tensor(self) -> torch.Tensor:
        """The buffer holding all tables, media and materials"""
        return self._tensor

    def flush(self) -> None:
        """Uploads the host copy. Does nothing if the store is frozen."""
        if not self.frozen:
            self._tensor[: self._host.size].copy_(torch.from_numpy(self._host))

    def _hostPtr(self, offset: int) -> int:
        return self._host.ctypes.data + offset

    def _updateTable(self, name: str, data) -> int:
        if data is None:
            return 0
        if name not in self._tableOff:
            raise ValueError(f"Table {name} has not been previously allocated")
        off, size = self._tableOff[name]
        if getTableSize(data) > size:
            raise ValueError(f"Table {name} does not fit in previous allocation")
        Table(data).copy(self._hostPtr(off))
        return self._tableAdr[name]

    def updateMedium(self, medium: Medium) -> None:
        if self.frozen:
            raise RuntimeError("Cannot update frozen MaterialStore")
        if medium.name not in self._mediaOff:
            raise ValueError(f"Medium {medium.name} has not been previously allocated")
        glsl = Medium.GLSL.from_address(self._hostPtr(self._mediaOff[medium.name]))
        glsl.lambda_min = medium.lambda_min
        glsl.lambda_max = medium.lambda_max
        for prop, key, field in _TABLES:
            setattr(glsl, field, self._updateTable(f"{medium.name}_{key}", getattr(medium, prop)))

    def updateMaterial(self, material: Material, *, updateMedia: bool = False) -> None:
        if self.frozen:
            raise RuntimeError("Cannot update frozen MaterialStore")
        if material.name not in self._matOff:
            raise ValueError(f"Material {material.name} has not been previously allocated")
        inside, outside = _mediumName(material.inside), _mediumName(material.outside)
        for ref in (inside, outside):
            if ref is not None and ref not in self.media:
                raise ValueError(f"Material {material.name} references unknown medium {ref}")
        glsl = Material.GLSL.from_address(self._hostPtr(self._matOff[material.name]))
        glsl.inside = 0 if inside is None else self.media[inside]
        glsl.outside = 0 if outside is None else self.media[outside]
        glsl.flagsInwards = int(material.flagsInward)
        glsl.flagsOutwards = int(material.flagsOutward)
        if updateMedia:
            for m in (material.inside, material.outside):
                if isinstance(m, Medium):
                    self.updateMedium(m)


# ------------------------------------------------------------------ persistence
def saveMaterials(path, material: Iterable[Material], *, media: Iterable[Medium] = []) -> None:
    """
    Stores materials and all media they reference in the archive layout of the reference
    (src/theia/material.py:715-794): ``material.json`` (name, inside, outside, flagsInward,
    flagsOutward) next to ``media/<name>`` (numpy ``.npz`` content, see ``Medium.save``), in a
    directory if ``path`` is one, else packed into a zip file.
    """
    material = list(material)
    mediaDict = {m.name: m for m in media}
    for mat in material:
        for m in (mat.inside, mat.outside):
            if isinstance(m, Medium):
                mediaDict.setdefault(m.name, m)
    for mat in material:
        for m in (mat.inside, mat.outside):
            if type(m) is str and m not in mediaDict:
                raise ValueError(f"Material '{mat.name}' references an unknown medium '{m}'!")
    entries = [
        {"name": mat.name, "inside": _mediumName(mat.inside), "outside": _mediumName(mat.outside),
         "flagsInward": int(mat.flagsInward), "flagsOutward": int(mat.flagsOutward)}
        for mat in material
    ]
    path = Path(path)
    if path.is_dir():
        (path / "media").mkdir(exist_ok=True)
        for name, m in mediaDict.items():
            with open(path / "media" / name, "wb") as f:
                m.save(f)
        (path / "material.json").write_text(json.dumps(entries))
    elif path.exists():
        raise ValueError("Invalid path!")
    else:
        with ZipFile(path, "w") as z:
            for name, m in mediaDict.items():
                with z.open(f"media/{name}", "w") as f:
                    m.save(f)
            z.writestr("material.json", json.dumps(entries))


def loadMaterials(path, *, skipValidation: bool = False) -> tuple[dict[str, Material], dict[str, Medium]]:
    """
    Inverse of ``saveMaterials`` (src/theia/material.py:797-881): reads an archive written by this
    package or by the reference and returns the materials and the media by name.
    """
    import io
    path = Path(path)
    media: dict[str, Medium] = {}
    if path.is_file():
        if not is_zipfile(path):
            raise ValueError("Unexpected file format!")
        with ZipFile(path) as z:
            names = [n for n in z.namelist() if n.startswith("media/") and not n.endswith("/")]
            if not names and not any(n.startswith("media") for n in z.namelist()):
                raise ValueError('Subdirectory "media" is missing!')
            for n in names:
                stem = Path(n).stem
                try:
                    media[stem] = Medium.load(io.BytesIO(z.read(n)), name=stem)
                except Exception:
                    raise ValueError(f"Failed to load media file: {n}")
            if "material.json" not in z.namelist():
                raise ValueError('Missing file: "material.json"!')
            raw = z.read("material.json").decode()
    elif path.is_dir():
        mediaDir = path / "media"
        if not mediaDir.is_dir():
            raise ValueError('Subdirectory "media" is missing!')
        for f in sorted(mediaDir.iterdir()):
            try:
                with open(f, "rb") as fh:
                    media[f.stem] = Medium.load(fh, name=f.stem)
            except Exception:
                raise ValueError(f"Failed to load media file: {f.absolute()}")
        if not (path / "material.json").is_file():
            raise ValueError('Missing file: "material.json"!')
        raw = (path / "material.json").read_text()
    else:
        raise ValueError("Incompatible path!")
    try:
        entries = json.loads(raw)
    except Exception:
        raise ValueError('Invalid material description file: "material.json"!')
    required = {"name", "inside", "outside", "flagsInward", "flagsOutward"}
    mats: dict[str, Material] = {}
    for e in entries:
        if not skipValidation and (not isinstance(e, dict) or set(e) != required):
            raise ValueError('Invalid material description file: "material.json"!')
        name = e["name"]
        if name in mats:
            raise ValueError(f"Duplicate material '{name}'!")
        sides = []
        for med in (e["inside"], e["outside"]):
            if med is not None and med not in media:
                raise ValueError(f"Material '{name}' references unknown medium: '{med}'!")
            sides.append(None if med is None else media[med])
        mats[name] = Material(name, sides[0], sides[1],
                              flags=(MaterialFlags(e["flagsInward"]), MaterialFlags(e["flagsOutward"])))
    return mats, media


# ------------------------------------------------------------------ models
def _dataFile(name: str):
    return importlib.resources.files("theia_b200").joinpath(f"data/{name}")


class MediumModel:
    """
    Base class of medium models: subclasses override the property functions they define and
    ``createMedium`` samples them into tables (src/theia/material.py:1123-1256).
    """

    ModelName = "noname"

    def refractive_index(self, wavelength):
        return None

    def group_velocity(self, wavelength):
        return None

    def absorption_coef(self, wavelength):
        return None

    def scattering_coef(self, wavelength):
        return None

    def log_phase_function(self, cos_theta):
        return None

    def phase_sampling(self, eta):
        return None

    def phase_m12(self, cos_theta):
        return None

    def phase_m22(self, cos_theta):
        return None

    def phase_m33(self, cos_theta):
        return None

    def phase_m34(self, cos_theta):
        return None

    def createMedium(self, lambda_min=200.0 * u.nm, lambda_max=800.0 * u.nm, num_lambda=1024, num_theta=1024,
                     *, name: str | None = None) -> Medium:
        """Samples the model: wavelengths on [lambda_min, lambda_max], cos(theta) on [-1, 1]."""
        lam = np.linspace(lambda_min, lambda_max, num_lambda)
        cos_theta = np.linspace(-1.0, 1.0, num_theta)
        eta = np.linspace(0.0, 1.0, num_theta)
        return Medium(
            name if name is not None else self.ModelName, lambda_min, lambda_max,
            refractive_index=self.refractive_index(lam),
            group_velocity=self.group_velocity(lam),
            absorption_coef=self.absorption_coef(lam),
            scattering_coef=self.scattering_coef(lam),
            log_phase_function=self.log_phase_function(cos_theta),
            phase_sampling=self.phase_sampling(eta),
            phase_m12=self.phase_m12(cos_theta),
            phase_m22=self.phase_m22(cos_theta),
            phase_m33=self.phase_m33(cos_theta),
            phase_m34=self.phase_m34(cos_theta),
        )


class SellmeierEquation:
    """n^2 = 1 + sum_i B_i L^2 / (L^2 - C_i), L in nm, C_i in nm^2 (src/theia/material.py:1259-1302)."""

    def __init__(self, B1: float, B2: float, B3: float, C1: float, C2: float, C3: float) -> None:
        self.B1, self.B2, self.B3 = B1, B2, B3
        self.C1, self.C2, self.C3 = C1, C2, C3

    def refractive_index(self, wavelength: ArrayLike) -> NDArray:
        L2 = np.square(u.convert(wavelength, u.nm))
        terms = [B * L2 / (L2 - C) for B, C in ((self.B1, self.C1), (self.B2, self.C2), (self.B3, self.C3))]
        return np.sqrt(1.0 + terms[0] + terms[1] + terms[2])

    def group_velocity(self, wavelength: ArrayLike) -> NDArray:
        """vg = c / (n - L dn/dL), with dn/dL = -sum_i B_i C_i L / (L^2 - C_i)^2 / n"""
        wavelength = np.asarray(wavelength)
        n = self.refractive_index(wavelength)
        L = u.convert(wavelength, u.nm)
        L2 = np.square(wavelength)
        dsum = sum(B * C * L / np.square(L2 - C) for B, C in ((self.B1, self.C1), (self.B2, self.C2), (self.B3, self.C3)))
        grad = -1.0 * dsum / n
        return 1.0 / (n - wavelength * grad) * u.c


class BK7Model(SellmeierEquation, MediumModel):
    """
    Schott N-BK7: Sellmeier refractive index and an absorption coefficient derived from the
    data sheet's internal transmittance at 10 and 25 mm (src/theia/material.py:1305-1358).
    """

    ModelName = "bk7"
    TransmissionTable = None

    def __init__(self) -> None:
        super().__init__(1.03961212, 0.231792344, 1.010469450, 0.00600069867e6, 0.0200179144e6, 103.5606530e6)
        if BK7Model.TransmissionTable is None:
            BK7Model.TransmissionTable = np.loadtxt(_dataFile("bk7_transmission.csv"), delimiter=",", skiprows=2)

    def absorption_coef(self, wavelength: ArrayLike) -> NDArray:
        # Beer-Lambert on both probes; average the absorption LENGTHS weighted by thickness
        # (so a measured transmission of zero contributes length 0 instead of an infinite mu)
        tab = BK7Model.TransmissionTable
        with np.errstate(divide="ignore"):
            tau10 = -0.010 / np.log(tab[:, 1])
            tau25 = -0.025 / np.log(tab[:, 2])
            tau = (10.0 * tau10 + 25.0 * tau25) / 35.0
            tau = np.interp(u.convert(np.asarray(wavelength), u.nm), tab[:, 0], tau)
            return np.reciprocal(tau) / u.m


class HenyeyGreensteinPhaseFunction:
    """Henyey-Greenstein phase function with asymmetry g (src/theia/material.py:1361-1419)."""

    def __init__(self, g: float = 0.0) -> None:
        self.g = g

    @property
    def g(self) -> float:
        return self._g

    @g.setter
    def g(self, value: float) -> None:
        self._g = value
        if not -1.0 < value < 1.0:
            warnings.warn("Asymmetry parameter outside the valid range (-1,1)!", RuntimeWarning)

    def log_phase_function(self, cos_theta: ArrayLike) -> NDArray:
        """log of (1-g^2) / (1+g^2-2g cos)^(3/2) / 4pi"""
        g = self.g
        cos_theta = np.asarray(cos_theta)
        return np.log((1.0 - g**2) / np.power(1.0 + g**2 - 2 * g * cos_theta, 1.5) / (4.0 * np.pi))

    def phase_sampling(self, eta: ArrayLike) -> NDArray:
        """inverse CDF: cos(theta) for unit random numbers eta"""
        g = self.g
        eta = np.asarray(eta)
        if abs(g) < 1e-7:
            return 1.0 - 2.0 * eta
        return (1.0 + g**2 - ((1.0 - g**2) / (1 + g - 2.0 * g * eta)) ** 2) / (2.0 * g)


class FournierForandPhaseFunction:
    """
    Fournier-Forand phase function for particles of refractive index n with a Junge size
    distribution of slope mu (src/theia/material.py:1422-1514).
    """

    def __init__(self, n: float, mu: float) -> None:
        self._n = n
        self._mu = mu
        self._update()

    @property
    def n(self) -> float:
        return self._n

    @n.setter
    def n(self, value: float) -> None:
        self._n = value
        self._update()

    @property
    def mu(self) -> float:
        return self._mu

    @mu.setter
    def mu(self, value: float) -> None:
        self._mu = value
        self._update()

    def _consts(self, cos_theta):
        nu = 0.5 * (3.0 - self.mu)
        k = 3.0 * (self.n - 1.0) ** 2
        d = 2.0 * (1.0 - cos_theta) / k
        d180 = 4.0 / k
        return nu, d, np.float_power(d, nu), d180, np.float_power(d180, nu)

    def log_phase_function(self, cos_theta: ArrayLike) -> NDArray:
        x = np.clip(cos_theta, -1.0, 1.0 - 1e-7)  # singular at cos = 1
        nu, d, d_nu, d180, d180_nu = self._consts(x)
        A = nu * (1 - d) - (1 - d_nu) + 2 * (d * (1 - d_nu) - nu * (1 - d)) / (1 - x)
        B = 4 * np.pi * (1 - d) ** 2 * d_nu
        C = (1 - d180_nu) * (3 * x**2 - 1)
        D = 16 * np.pi * (d180 - 1) * d180_nu
        return np.log(A / B + C / D)

    def phase_sampling(self, eta: ArrayLike) -> NDArray:
        return self._sample_spline(eta)

    def _update(self) -> None:
        # tabulate the analytic CDF and interpolate its inverse with a cubic spline
        from scipy.interpolate import CubicSpline

        x = np.linspace(1.0 - 1e-7, -1.0, 2048)
        _, d, d_nu, d180, d180_nu = self._consts(x)
        A = ((1 - d_nu * d) - 0.5 * (1 - d_nu) * (1 - x)) / ((1 - d) * d_nu)
        B = ((1 - d180_nu) * (1 - x) * x) / (16 * (d180 - 1) * d180_nu)
        self._sample_spline = CubicSpline(A + B, x)


class DispersionFreeMedium:
    """Medium with wavelength independent n, group index, mu_a, mu_s (src/theia/material.py:1517-1593)."""

    def __init__(self, *, n: float = 1.0, ng: float = 1.0, mu_a: float = 0.0, mu_s: float = 0.0) -> None:
        self.n = n
        self.ng = ng
        self.mu_a = mu_a
        self.mu_s = mu_s

    # Properties backed by _n, _ng, _mu_a, _mu_s like the reference's: a model that also mixes in
    # FournierForandPhaseFunction shares `_n` with it there, and is reproduced as is.
    n = property(lambda self: self._n, lambda self, v: setattr(self, "_n", v), doc="Refractive index")
    ng = property(lambda self: self._ng, lambda self, v: setattr(self, "_ng", v), doc="Group index")
    mu_a = property(lambda self: self._mu_a, lambda self, v: setattr(self, "_mu_a", v), doc="Absorption coefficient")
    mu_s = property(lambda self: self._mu_s, lambda self, v: setattr(self, "_mu_s", v), doc="Scattering coefficient")

    def refractive_index(self, wavelength):
        return np.ones_like(wavelength) * self.n

    def group_velocity(self, wavelength):
        return np.ones_like(wavelength) / self.ng * u.c

    def absorption_coef(self, wavelength):
        return np.ones_like(wavelength) * self.mu_a

    def scattering_coef(self, wavelength):
        return np.ones_like(wavelength) * self.mu_s


class WaterBaseModel:
    """
    (Sea) water without phase function: refractive index after Millard & Seaver (1990) as a
    function of temperature [C], pressure [dbar] and salinity [PSU]; absorption and
    scattering interpolated from Smith & Baker (1981) (src/theia/material.py:1596-1790).
    """

    DataTable = None
    # coefficients of the Millard & Seaver 27 term fit, keyed by (power of L, T, p, S);
    # L^-2 etc. are written as negative powers
    A0 = 1.3280657
    L2 = -0.0045536802
    LM2 = 0.0025471707
    LM4 = 0.000007501966
    LM6 = 0.000002802632
    T1 = -0.0000052883907
    T2 = -0.0000030738272
    T3 = 0.000000030124687
    T4 = -2.0863178e-10
    TL = 0.000010508621
    T2L = 0.00000021282248
    T3L = -0.000000001705881
    S0 = 0.00019029121
    S1LM2 = 0.0000024239607
    S1T = -0.00000073960297
    S1T2 = 0.0000000089818478
    S1T3 = 1.2078804e-10
    STL = -0.0000003589495
    P1 = 0.0000015868363
    P2 = -1.574074e-11
    PLM2 = 0.000000010712063
    PT = -0.0000000094634486
    PT2 = 1.0100326e-10
    P2T2 = 5.8085198e-15
    P1S = -0.0000000011177517
    PTS = 5.7311268e-11
    PT2S = -1.5460458e-12

    def __init__(self, temperature: float, pressure: float, salinity: float) -> None:
        self.temperature = temperature
        self.pressure = pressure
        self.salinity = salinity
        if WaterBaseModel.DataTable is None:
            WaterBaseModel.DataTable = np.loadtxt(_dataFile("water_smith81.csv"), delimiter=",", skiprows=1)

    @property
    def temperature(self) -> float:
        return self._T

    @temperature.setter
    def temperature(self, value: float) -> None:
        self._T = value
        if not 0.0 <= value <= 30.0:
            warnings.warn("Temperature is outside the models valid range of 0°-30°C", RuntimeWarning)

    @property
    def pressure(self) -> float:
        return self._p

    @pressure.setter
    def pressure(self, value: float) -> None:
        self._p = value
        if not 0.0 <= value <= 11_000:
            warnings.warn("Pressure is outside the models valid range of 0-11.000 dbar", RuntimeWarning)

    @property
    def salinity(self) -> float:
        return self._S

    @salinity.setter
    def salinity(self, value: float) -> None:
        self._S = value
        if not 0.0 <= value <= 40.0:
            warnings.warn("Salinity is outside the models valid range of 0-40 psu", RuntimeWarning)

    def refractive_index(self, wavelength: ArrayLike) -> NDArray:
        L = u.convert(wavelength, u.um)
        T, p, S = self.temperature, self.pressure, self.salinity
        N1 = (self.A0 + self.L2 * (L**2) + self.LM2 / (L**2) + self.LM4 / (L**4) + self.LM6 / (L**6)
              + self.T1 * T + self.T2 * (T**2) + self.T3 * (T**3) + self.T4 * (T**4)
              + self.TL * T * L + self.T2L * (T**2) * L + self.T3L * (T**3) * L)
        N2 = (self.S0 * S + self.S1LM2 * S / (L**2) + self.S1T * S * T + self.S1T2 * S * (T**2)
              + self.S1T3 * S * (T**3) + self.STL * S * T * L)
        N3 = (self.P1 * p + self.P2 * (p**2) + self.PLM2 * p / (L**2) + self.PT * p * T
              + self.PT2 * p * (T**2) + self.P2T2 * (p**2) * (T**2))
        N4 = self.P1S * p * S + self.PTS * p * T * S + self.PT2S * p * (T**2) * S
        return N1 + N2 + N3 + N4

    def group_velocity(self, wavelength: ArrayLike) -> NDArray:
        """vg = c / (n - L dn/dL), derivative of the fit taken term by term"""
        L = u.convert(wavelength, u.um)
        T, p, S = self.temperature, self.pressure, self.salinity
        G1 = (2.0 * self.L2 * L - 2.0 * self.LM2 / (L**3) - 4.0 * self.LM4 / (L**5) - 6.0 * self.LM6 / (L**7)
              + self.TL * T + self.T2L * (T**2) + self.T3L * (T**3))
        G2 = -2.0 * self.S1LM2 * S / (L**3) + self.STL * S * T
        G3 = -2.0 * self.PLM2 * p / (L**3)
        G = G1 + G2 + G3 + 0.0
        n = self.refractive_index(wavelength)
        return 1.0 / (n - L * G) * u.c

    def absorption_coef(self, wavelength: ArrayLike) -> NDArray:
        tab = WaterBaseModel.DataTable
        return np.interp(u.convert(wavelength, u.nm), tab[:, 0], tab[:, 1]) / u.m

    def scattering_coef(self, wavelength: ArrayLike) -> NDArray:
        tab = WaterBaseModel.DataTable
        return np.interp(u.convert(wavelength, u.nm), tab[:, 0], tab[:, 2]) / u.m


class KokhanovskyOceanWaterPhaseMatrix:
    """
    Kokhanovsky's parameterisation of the reduced Mueller matrix of oceanic water
    (JGR 108 C6, 2003): p90 degree of polarisation at 90 deg, theta0 angle shift, alpha and
    xi slope and scale of the multiple scattering term (src/theia/material.py:1793-1878).
    """

    def __init__(self, p90, theta0, alpha, xi) -> None:
        self.p90 = p90
        self.theta0 = theta0
        self.alpha = alpha
        self.xi = xi

    def phase_m12(self, cos_theta: ArrayLike) -> NDArray:
        c2 = np.square(cos_theta)
        return -self.p90 * (1.0 - c2) / (1.0 + self.p90 * c2)

    def phase_m22(self, cos_theta: ArrayLike) -> NDArray:
        theta = np.arccos(cos_theta)
        cz2 = np.square(np.cos(theta - self.theta0))
        e = self.xi * np.exp(-self.alpha * theta)
        return (self.p90 * (1.0 + cz2) + e) / (1.0 + self.p90 * cz2 + e)

    def phase_m33(self, cos_theta: ArrayLike) -> NDArray:
        cos_theta = np.asarray(cos_theta)
        theta = np.arccos(cos_theta)
        e = self.xi * np.exp(-self.alpha * theta)
        return (2 * self.p90 * cos_theta + e) / (1.0 + self.p90 * np.square(cos_theta) + e)
