"""
Digests of the hits of the parity scenarios (tests/golden/oracle_digests.json).

The parity tests compare the CUDA path with the oracle run side by side; both could drift together (a change of the
arithmetic contract made on both sides) and stay equal. The digests pin today's results: SHA-256 over the hits of
every scenario in canonical order - sorted by (path, event tag), then path, tag, objectId and every float field as
little endian float32 bytes. They are outputs of THIS repo's oracle (tests/golden/generate_oracle_digests.py), not of
the reference, whose tracers cannot run here: a regression anchor, not a parity claim.
"""

from __future__ import annotations

import hashlib
import json
import os

import numpy as np

import harness as H

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "oracle_digests.json")
N_PATHS = 1 << 12

# name -> (builder, keyword arguments); response is always the hit recorder
CASES = {
    "C1": (H.volumeScenario, dict()),
    "C1,mis=False,polarized,water": (H.volumeScenario, dict(mis=False, polarized=True, water=True)),
    "C1,direct=False,cone": (H.volumeScenario, dict(direct=False, lightKind="cone")),
    "C1,flat target": (H.volumeScenario, dict(targetKind="flat")),
    "C2": (H.sceneC2, dict()),
    "C2,cone": (H.sceneC2, dict(lightKind="cone")),
    "C2,unpolarized,refracted hit dir": (H.sceneC2, dict(polarized=False, lightKind="cone", useRefractedHitDir=True)),
    "C3": (H.sceneC3, dict()),
    "C3,no guide": (H.sceneC3, dict(guide=False)),
    "C3,polarized,no direct": (H.sceneC3, dict(polarized=True, disableDirectLighting=True)),
    "C3,small guide": (H.sceneC3, dict(guide="small")),
    "C5,6x6x6": (H.sceneC5, dict(resp="record", lattice=(6, 6, 6), pitch=(4.0, 4.0, 4.0))),
}


def build(name: str, device: str):
    from theia_b200 import _device
    builder, kw = CASES[name]
    kw = dict(kw)
    kw.setdefault("resp", "record")
    with _device.useDevice(device):
        return builder(N_PATHS, **kw)


def digestOf(hits: dict) -> dict:
    """hits: harness.sortedHits(...)"""
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(hits["path"], dtype="<i8").tobytes())
    h.update(np.ascontiguousarray(hits["tag"], dtype="<i8").tobytes())
    h.update(np.ascontiguousarray(hits["objectId"], dtype="<i4").tobytes())
    for f in H.FIELDS + H.POL_FIELDS:
        if f == "objectId" or f not in hits:
            continue
        h.update(np.ascontiguousarray(hits[f], dtype="<f4").tobytes())
    return {"count": int(hits["count"]), "sha256": h.hexdigest(),
            "contrib_sum": float(np.sum(hits["contrib"], dtype=np.float64))}


def load() -> dict:
    with open(PATH) as f:
        return json.load(f)
