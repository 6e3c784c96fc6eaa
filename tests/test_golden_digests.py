"""
The committed digests of the parity scenarios (tests/golden/oracle_digests.json, see tests/digests.py): the oracle
reproduces them (CPU suite), and the CUDA path produces them on its own, with no oracle in the test (GPU suite). A
change that moves results on both sides at once shows up here.
"""

import pytest

import digests as D
import harness as H

GOLD = D.load()


def test_every_case_has_a_digest():
    assert GOLD["n_paths"] == D.N_PATHS
    assert sorted(GOLD["cases"]) == sorted(D.CASES)
    for name, d in GOLD["cases"].items():
        assert d["count"] > 30 and len(d["sha256"]) == 64, name


@pytest.mark.parametrize("name", list(D.CASES))
def test_oracle_reproduces_the_digest(orc, name):
    cfg = D.build(name, "cpu")
    H.runOracle(cfg)
    assert D.digestOf(H.sortedHits(cfg)) == GOLD["cases"][name]


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["single", "wavefront"])
@pytest.mark.parametrize("name", list(D.CASES))
def test_cuda_path_produces_the_digest(cuda, name, mode):
    builder, kw = D.CASES[name]
    if builder is H.volumeScenario:
        if mode != "single":
            pytest.skip("the volume tracer has one kernel")
        cfg = D.build(name, "cuda")
    else:
        from theia_b200 import _device
        with _device.useDevice("cuda"):
            cfg = builder(D.N_PATHS, **dict(kw, resp="record", mode=mode))
    H.runGpu(cfg)
    assert D.digestOf(H.sortedHits(cfg)) == GOLD["cases"][name]
