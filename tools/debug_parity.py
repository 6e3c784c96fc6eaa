"""Prints the first hits that differ between the CUDA path and the oracle for a harness scenario.
    python tools/debug_parity.py volumeScenario targetKind=flat
"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H
from theia_b200 import _device

def parse(v):
    if v in ("True", "False"): return v == "True"
    try: return int(v)
    except ValueError:
        try: return float(v)
        except ValueError: return v

builder = getattr(H, sys.argv[1])
kw = {a.split("=")[0]: parse(a.split("=")[1]) for a in sys.argv[2:]}
N = kw.pop("N", 1 << 14)
kw.setdefault("resp", "record")
with _device.useDevice("cuda"): g = builder(N, **kw)
with _device.useDevice("cpu"): c = builder(N, **kw)
H.runGpu(g); H.runOracle(c)
a, b = H.sortedHits(g), H.sortedHits(c)
print("counts", a["count"], b["count"])
ka = {(p, t): i for i, (p, t) in enumerate(zip(a["path"], a["tag"]))}
kb = {(p, t): i for i, (p, t) in enumerate(zip(b["path"], b["tag"]))}
onlyA = sorted(set(ka) - set(kb))[:5]; onlyB = sorted(set(kb) - set(ka))[:5]
print("only cuda", onlyA); print("only oracle", onlyB)
for key in onlyA[:3]:
    i = ka[key]; print("cuda", key, {f: a[f][i] for f in H.FIELDS})
for key in onlyB[:3]:
    i = kb[key]; print("orc ", key, {f: b[f][i] for f in H.FIELDS})
paths = sorted({k[0] for k in onlyA + onlyB})[:3]
for p in paths:
    for name, d in (("cuda", a), ("orc", b)):
        m = d["path"] == p
        print(name, "path", p, "tags", d["tag"][m], "contrib", d["contrib"][m], "time", d["time"][m])
common = sorted(set(ka) & set(kb))
bad = 0
for key in common:
    i, j = ka[key], kb[key]
    for f in H.FIELDS:
        if not np.array_equal(a[f][i], b[f][j], equal_nan=True):
            print("value differs", key, f, a[f][i], b[f][j]); bad += 1; break
    if bad >= 5: break
