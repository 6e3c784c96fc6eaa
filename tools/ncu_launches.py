"""
Summarises an ncu launch list (long CSV of `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
--csv --log-file ...` over a bench.py run) per kernel class of the wavefront scene tracer: launches, time, share of a
step, DRAM bytes per launch.

    python tools/ncu_launches.py gpurun_out/r2u_launches.csv c3 [--traffic profiles/traffic.json] > profiles/r2_launches_c3.txt

Only complete steps (generate ... next generate) are counted, the first one (module load, cold caches) is left out.
The ray query launches of a step alternate primary / shadow (target guide) and are classed by their position. With
--traffic the DRAM bytes per launch of every class are written where bench.py looks for `roofline.traffic`.
"""

from __future__ import annotations

import collections
import csv
import json
import re
import sys


def short(name: str) -> str:
    m = re.search(r"(wf\w+Kernel|histReduceKernel|volumeForwardKernel|sceneForwardKernel|\w+Kernel)", name)
    if "at::" in name or "vectorized_elementwise" in name:
        return "torch elementwise"
    return m.group(1) if m else name[:40]


def main() -> None:
    path, workload = sys.argv[1], sys.argv[2]
    traffic = sys.argv[sys.argv.index("--traffic") + 1] if "--traffic" in sys.argv else None
    launches = collections.OrderedDict()
    for r in csv.reader(open(path)):
        if len(r) < 15 or not r[0].isdigit():
            continue
        d = launches.setdefault(int(r[0]), {"name": short(r[4]), "grid": r[8]})
        v = float(r[14].replace(",", ""))
        unit = r[13]
        if r[12] == "gpu__time_duration.sum":
            d["us"] = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)
        else:
            d[r[12]] = v * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
    seq = list(launches.values())
    # steps: from one generate kernel to the next
    starts = [i for i, d in enumerate(seq) if d["name"] == "wfGenerateKernel"]
    steps = [seq[a:b] for a, b in zip(starts[:-1], starts[1:])][1:]
    if not steps:
        raise SystemExit("no complete step in the launch list")
    agg = collections.OrderedDict()
    for st in steps:
        guide = any(d["name"] == "wfShadeShadowKernel" for d in st)
        nq = 0
        for d in st:
            cls = d["name"]
            if cls in ("wfTraceQueuedKernel", "wfTraceKernel", "wfTraceSimpleKernel"):
                # generate, then per generation: trace, shade, [shadow trace, shadow shade]
                prevShade = nq > 0 and lastName == "wfShadeKernel"  # noqa: F821
                cls = cls + ("(shadow)" if guide and prevShade else "")
                nq += 1
            lastName = d["name"]  # noqa: F841
            a = agg.setdefault(cls, {"n": 0, "us": 0.0, "rd": 0.0, "wr": 0.0})
            a["n"] += 1
            a["us"] += d.get("us", 0.0)
            a["rd"] += d.get("dram__bytes_read.sum", 0.0)
            a["wr"] += d.get("dram__bytes_write.sum", 0.0)
    total = sum(a["us"] for a in agg.values())
    print(f"# {path}: {len(steps)} complete steps of workload {workload} (first step left out); times under ncu are serialised,")
    print("# cold-cache per-launch durations: the SHARE of a step is what compares with bench.py's live event times")
    print(f"{'kernel class':34s} {'launches/step':>13s} {'us/launch':>10s} {'share':>7s} {'DRAM rd MB/launch':>18s} {'DRAM wr MB/launch':>18s}")
    out = {}
    for cls, a in sorted(agg.items(), key=lambda x: -x[1]["us"]):
        print(f"{cls:34s} {a['n'] / len(steps):13.1f} {a['us'] / a['n']:10.1f} {a['us'] / total:7.3f} "
              f"{a['rd'] / a['n'] / 1e6:18.2f} {a['wr'] / a['n'] / 1e6:18.2f}")
        out[cls] = {"launch_class": "average over one step", "dram_bytes_per_launch": (a["rd"] + a["wr"]) / a["n"],
                    "launches_per_step": a["n"] / len(steps), "source": path}
    print(f"{'step total (ms)':34s} {total / len(steps) / 1e3:24.3f}")
    if traffic:
        try:
            cur = json.load(open(traffic))
        except Exception:
            cur = {}
        # bench.py names the ray query classes after the phased kernel whatever kernel ran
        ren = {"wfTraceQueuedKernel": "wfTraceKernel", "wfTraceSimpleKernel": "wfTraceKernel",
               "wfTraceQueuedKernel(shadow)": "wfTraceKernel(shadow)", "wfTraceSimpleKernel(shadow)": "wfTraceKernel(shadow)"}
        cur[workload] = {ren.get(k, k): v for k, v in out.items()}
        json.dump(cur, open(traffic, "w"), indent=1)


if __name__ == "__main__":
    main()
