"""Summarises an ncu launch list (--metrics gpu__time_duration.sum --csv): time per kernel.
    python tools/launch_summary.py gpurun_out/launches.csv [skip_launches]
"""
import collections
import csv
import sys

rows = [r for r in csv.DictReader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
t = collections.defaultdict(lambda: [0, 0.0])
for r in rows[skip:]:
    k = r["Kernel Name"].split("(")[0][:48]
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
    t[k][0] += 1
    t[k][1] += v
tot = sum(v[1] for v in t.values())
print(f"{len(rows) - skip} launches, {tot / 1e3:.3f} ms")
for k, v in sorted(t.items(), key=lambda x: -x[1][1]):
    print(f"{k:50s} n={v[0]:5d} total_us={v[1]:11.1f} share={v[1] / tot:.3f}")
