"""Runs bench.py for several workloads / env settings and prints one compact line each.
    python tools/quick.py c2,c3,c5 [KEY=V1,V2 ...] [-- extra bench args]
"""
import itertools
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = sys.argv[1:]
extra = []
if "--" in args:
    k = args.index("--")
    args, extra = args[:k], args[k + 1:]
workloads = args[0].split(",")
sweeps = [(a.split("=")[0], a.split("=")[1].split(",")) for a in args[1:]]
for combo in itertools.product(*[v for _, v in sweeps]) if sweeps else [()]:
    env = dict(os.environ)
    env.update({k: v for (k, _), v in zip(sweeps, combo)})
    for w in workloads:
        cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--workload", w, "--steps", "5", "--warmup", "3", "--no-cpu",
               "--secondary", "", "--e2e-small-batch", "0", *extra]
        try:
            r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=240)
            d = json.loads(r.stdout.strip().splitlines()[-1])
            print(" ".join(f"{k}={v}" for (k, _), v in zip(sweeps, combo)), w, f"{d['value']:.4g} seg/s", f"{d['ms_per_step']:.3f} ms/step",
                  f"e2e {d['e2e']['value']:.4g}",
                  " ".join(f"{n.replace('Kernel', '')}={v['ms_per_launch']:.3f}" for n, v in d["roofline"]["kernels"].items()), flush=True)
        except Exception as e:  # noqa: BLE001
            print(combo, w, "FAILED", repr(e)[:200], flush=True)
