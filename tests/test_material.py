"""
The medium models that fill the tables the tracers sample, against the reference's golden
vectors (tests/golden/*.csv are the reference's tests/data/*.csv; the assertions follow the
reference's tests/test_material.py:11-127), plus the table layout the oracle and the CUDA
kernels consume (sh/lookup.glsl:14-29, sh/material.glsl:11-74).
"""

import os
import warnings

import numpy as np
from scipy.integrate import quad

import theia_b200.units as u
from theia_b200 import _device, material

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
DATA = os.path.join(os.path.dirname(HERE), "theia_b200", "data")


def numericGroupVelocity(model):
    l = np.linspace(300.0, 800.0, 200) * u.nm
    n = model.refractive_index(l)
    return l, 1.0 / (n - l * np.gradient(n, 500 / (len(l) - 1))) * u.c


def test_bk7_against_schott_table():
    model = material.BK7Model()
    ref = np.loadtxt(os.path.join(GOLDEN, "bk7_refractive_index.csv"), delimiter=",", skiprows=2)
    assert np.abs(ref[:, 1] - model.refractive_index(ref[:, 0] * u.nm)).max() < 5e-5
    l, vg = numericGroupVelocity(model)
    assert np.abs((vg - model.group_velocity(l)) / vg).max() < 1e-3
    trans = np.loadtxt(os.path.join(DATA, "bk7_transmission.csv"), delimiter=",", skiprows=2)
    mu_a = model.absorption_coef(trans[:, 0] * u.nm)
    t10, t25 = np.exp(-mu_a * 10.0 * u.mm), np.exp(-mu_a * 25.0 * u.mm)
    assert np.abs(t10 - trans[:, 1]).max() < 0.1 and np.abs(t10 - trans[:, 1]).mean() < 0.02
    assert np.abs(t25 - trans[:, 2]).max() < 0.01 and np.abs(t25 - trans[:, 2]).mean() < 2e-3


def test_water_against_austin_halikas_and_smith_baker():
    model = material.WaterBaseModel(10.0, 0.0, 35.0)
    ref = np.loadtxt(os.path.join(GOLDEN, "water_n_10C_35S.csv"), delimiter=",", skiprows=3)
    assert np.abs(ref[:, 1] - model.refractive_index(ref[:, 0] * u.nm)).max() < 0.005
    l, vg = numericGroupVelocity(model)
    assert np.abs((vg - model.group_velocity(l)) / vg).max() < 5e-3
    tab = np.loadtxt(os.path.join(DATA, "water_smith81.csv"), delimiter=",", skiprows=2)
    assert np.abs(tab[:, 1] - model.absorption_coef(tab[:, 0] * u.nm)).max() < 1e-6
    assert np.abs(tab[:, 2] - model.scattering_coef(tab[:, 0] * u.nm)).max() < 1e-6


def samplingError(model, bins=50, N=int(4e6)):  # 4x the reference sample: keeps 5e-4 at 3 sigma
    eta = np.random.default_rng(0xC0FFEE).random(N)
    h, edges = np.histogram(model.phase_sampling(eta), bins=bins)
    f = lambda x: np.exp(model.log_phase_function(x))  # noqa: E731
    want = [quad(f, edges[i], edges[i + 1])[0] * 2 * np.pi for i in range(bins)]
    return np.abs(h / N - want).max()


def phaseIntegral(model):
    return quad(lambda x: np.exp(model.log_phase_function(x)), -1.0, 1.0)[0] * 2 * np.pi


def test_henyey_greenstein_golden():
    ref = np.loadtxt(os.path.join(GOLDEN, "log_phase_hg.csv"), delimiter=",", skiprows=1)
    for col, g in enumerate((0.3, 0.0, -0.5, 0.99), start=1):
        hg = material.HenyeyGreensteinPhaseFunction(g)
        assert np.abs(ref[:, col] - hg.log_phase_function(ref[:, 0])).max() < 1e-6
        assert samplingError(hg) < 5e-4
        assert abs(phaseIntegral(hg) - 1.0) < 1e-5


def test_fournier_forand_golden():
    model = material.FournierForandPhaseFunction(1.175, 4.065)
    ref = np.loadtxt(os.path.join(GOLDEN, "log_phase_ff.csv"), delimiter=",", skiprows=1)
    assert np.all(np.abs(ref[:, 1] - model.log_phase_function(ref[:, 0])) < 1e-6)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        assert samplingError(model, 70, int(2e6)) < 0.01
        assert abs(phaseIntegral(model) - 1.0) < 1e-5


def test_medium_tables_and_store_layout():
    # material.py:1213-1256 of the reference: 1024 samples per table, lambda in [200, 800] nm,
    # angular tables over t = (cos + 1) / 2; the store packs {int nx; float samples[nx]}
    from harness import EnergyModel
    model = EnergyModel(0.05, 0.01, 0.3)
    med = model.createMedium()
    assert med.lambda_min == 200.0 and med.lambda_max == 800.0
    assert len(med.refractive_index) == 1024 and len(med.log_phase_function) == 1024
    cos = np.linspace(-1.0, 1.0, 1024)
    assert np.allclose(med.log_phase_function, model.log_phase_function(cos), atol=1e-6)
    assert np.allclose(med.phase_sampling, model.phase_sampling(np.linspace(0.0, 1.0, 1024)), atol=1e-6)
    with _device.useDevice("cpu"):
        store = material.MaterialStore([material.Material("m", med, None, flags=("TR", "B"))], media=[med])
        raw = store.tensor.numpy().view(np.uint8)
        base = store.tensor.data_ptr()
        off = store.media[med.name] - base
        words = raw[off:off + 88].view(np.uint64)  # 2 floats + 10 table pointers (sh/material.glsl:11-34)
        lam = raw[off:off + 8].view(np.float32)
        assert lam[0] == 200.0 and lam[1] == 800.0
        nPtr = int(words[1]) - base
        assert raw[nPtr:nPtr + 4].view(np.int32)[0] == 1024
        assert np.allclose(raw[nPtr + 4:nPtr + 4 + 4096].view(np.float32), 1.33)
        mat = store.material["m"] - base
        rec = raw[mat:mat + 24]
        assert rec[:8].view(np.uint64)[0] == store.media[med.name] and rec[8:16].view(np.uint64)[0] == 0
        fl = rec[16:24].view(np.uint32)
        assert fl[0] == int(material.parseMaterialFlags("TR")) and fl[1] == int(material.parseMaterialFlags("B"))


def test_material_archive_round_trip(tmp_path):
    # src/theia/material.py:715-881: material.json + media/<name> (npz), as a directory or a zip file
    import json
    from zipfile import ZipFile
    import pytest
    from theia_b200 import material
    from theia_b200.testing import WaterTestModel
    water = WaterTestModel().createMedium()
    glass = material.BK7Model().createMedium()
    mats = [material.Material("lens", glass, water, flags=("TR", "T")), material.Material("det", None, water, flags="DB")]
    for target in (tmp_path / "mats.zip", tmp_path / "dir"):
        if target.suffix != ".zip":
            target.mkdir()
        material.saveMaterials(target, mats)
        if target.suffix == ".zip":
            with ZipFile(target) as z:
                assert sorted(z.namelist()) == ["material.json", f"media/{glass.name}", f"media/{water.name}"]
                assert {e["name"] for e in json.loads(z.read("material.json"))} == {"lens", "det"}
        else:
            assert (target / "material.json").is_file() and (target / "media" / water.name).is_file()
        got, media = material.loadMaterials(target)
        assert set(got) == {"lens", "det"} and set(media) == {water.name, glass.name}
        assert got["lens"].flagsInward == mats[0].flagsInward and got["lens"].flagsOutward == mats[0].flagsOutward
        assert got["det"].inside is None and got["det"].outside.name == water.name
        for name, m in ((water.name, water), (glass.name, glass)):
            assert media[name].lambda_min == m.lambda_min and media[name].lambda_max == m.lambda_max
            for prop in material.Medium._PROPS:
                a, b = getattr(media[name], prop), getattr(m, prop)
                assert (a is None) == (b is None) and (a is None or np.array_equal(a, b))
        # the loaded materials build the same device tables
        s1, s2 = material.MaterialStore(mats), material.MaterialStore(list(got.values()))
        assert bytes(s1._tensor.cpu().numpy()[:64]) != b"" and s1._tensor.numel() == s2._tensor.numel()
    with pytest.raises(ValueError):
        material.saveMaterials(tmp_path / "bad.zip", [material.Material("m", "nowhere", None, flags="T")])
    with pytest.raises(ValueError):
        material.loadMaterials(tmp_path / "missing")
