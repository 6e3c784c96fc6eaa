"""
Attributes the per-instruction metrics of an ncu report (source page, SASS view) to CUDA source
lines, using the line table of the cubin inside libtheia_b200.so (nvdisasm -g).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep sceneForwardKernelILb1 [--top 40]

The .so must be the build the report was taken from (instruction order is matched 1:1).
"""

from __future__ import annotations

import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "theia_b200", "libtheia_b200.so")


def lineTable(kernel: str):
    """[(file:line, opcode)] of the kernel's instructions in address order"""
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", SO], cwd=d, check=True, capture_output=True)
        out = []
        for f in sorted(os.listdir(d)):
            if not f.endswith(".cubin"):
                continue
            txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, f)], capture_output=True, text=True).stdout
            inside, cur = False, "?"
            for ln in txt.splitlines():
                if ln.startswith("//---") and ".text." in ln:
                    inside = kernel in ln
                    cur = "?"
                    continue
                if not inside:
                    continue
                m = re.match(r'\s*//## File "(.*)", line (\d+)', ln)
                if m:
                    if "inlined at" not in ln:
                        cur = f"{os.path.basename(m.group(1))}:{m.group(2)}"
                    continue
                m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*?);", ln)
                if m:
                    out.append((cur, m.group(2).strip()))
            if out:
                return out
    raise SystemExit(f"kernel {kernel} not found")


def sassRows(report: str, kernel: str):
    txt = subprocess.run(["ncu", "-i", report, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    # find the block of the kernel
    blocks, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            blocks.append(cur)
        elif cur is not None:
            cur["rows"].append(r)
    for b in blocks:
        if kernel.replace("ILb1", "<(bool)1>").replace("ILb0", "<(bool)0>").split("<")[0].split("IL")[0] in b["name"]:
            hdr = b["rows"][0]
            return hdr, b["rows"][1:]
    raise SystemExit("kernel not in report")


def main() -> None:
    report, kernel = sys.argv[1], sys.argv[2]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 40
    table = lineTable(kernel)
    hdr, rows = sassRows(report, kernel)
    rows = [r for r in rows if r and r[0].startswith("0x")]
    if len(rows) != len(table):
        print(f"warning: {len(rows)} profiled instructions vs {len(table)} in the cubin", file=sys.stderr)
    ci = {h: i for i, h in enumerate(hdr)}
    want = ["# Samples", "Instructions Executed", "Thread Instructions Executed", "stall_long_sb", "stall_short_sb",
            "stall_wait", "stall_branch_resolving", "stall_no_inst", "stall_lg", "stall_math", "stall_barrier"]
    agg = collections.defaultdict(lambda: collections.Counter())
    tot = collections.Counter()
    for (loc, _op), r in zip(table, rows):
        for w in want:
            if w in ci:
                try:
                    v = float(r[ci[w]])
                except ValueError:
                    v = 0.0
                agg[loc][w] += v
                tot[w] += v
    byFile = collections.defaultdict(lambda: collections.Counter())
    for loc, c in agg.items():
        byFile[loc.split(":")[0]].update(c)
    print(f"total warp instr {tot['Instructions Executed']:.3e}  thread instr {tot['Thread Instructions Executed']:.3e}  "
          f"samples {tot['# Samples']:.0f}")
    print("\n== by file (share of samples | share of warp instr | avg active threads)")
    for f, c in sorted(byFile.items(), key=lambda x: -x[1]["# Samples"]):
        ie = c["Instructions Executed"]
        print(f"{f:28s} {c['# Samples'] / max(tot['# Samples'], 1):6.3f} {ie / max(tot['Instructions Executed'], 1):6.3f} "
              f"{c['Thread Instructions Executed'] / max(ie, 1):5.1f}")
    print(f"\n== top {top} lines by samples: loc samples% instr% threads/instr | dominant stalls")
    for loc, c in sorted(agg.items(), key=lambda x: -x[1]["# Samples"])[:top]:
        ie = c["Instructions Executed"]
        stalls = sorted(((k, v) for k, v in c.items() if k.startswith("stall_")), key=lambda x: -x[1])[:3]
        st = " ".join(f"{k[6:]}={v / max(c['# Samples'], 1):.2f}" for k, v in stalls if v > 0)
        print(f"{loc:28s} {100 * c['# Samples'] / max(tot['# Samples'], 1):5.1f} "
              f"{100 * ie / max(tot['Instructions Executed'], 1):5.1f} {c['Thread Instructions Executed'] / max(ie, 1):5.1f} | {st}")


if __name__ == "__main__":
    main()
