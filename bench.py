"""
bench.py - photon path segments/s of the forward tracing hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2] [--batch P]
    python bench.py --impl reference ...      # CPU arm: the oracle on the host cores

A step = one batch of P paths through the tracer kernel + the histogram reduce kernel.
N > 1: launched by torchrun, one rank per GPU; every rank traces its own batches (disjoint
Philox counter ranges, batch b = step * world + rank), no data-path collective, and ONE NCCL
all-reduce of the accumulated histogram at the end of the run (inside the timed region).

Timing: CUDA events on the launching stream around each step, L2 flushed (untimed) between
steps, max over ranks. "value" has all inputs resident in HBM (scene, media, parameters in
kernel argument space); "e2e" goes through the public pipeline API (stage updates, launch
assembly, reduce, device->host copy of the light curve into pinned memory) by wall clock.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "photon path segments/s"
UNIT = "segments/s"

# declared algorithmic HBM bytes per unit of each kernel class: state words
# (19 unpolarised / 26 polarised incl. the path number; dense buffers, no queue of slots) + ray query
# (32 B) + result (16 B)
def algorithmicBytes(polarized: bool) -> dict:
    state = 4 * (26 if polarized else 19)
    return {
        "volumeForwardKernel": 0,                                  # per segment: register resident
        "sceneForwardKernel": 0,
        "wfGenerateKernel": state + 32,                           # per path: write state, ray
        "wfTraceKernel": 32 + 16,                                 # per ray: read ray, write result
        "wfShadeKernel": 2 * state + 2 * 32 + 16,                  # per segment: read+write state/ray, read result
        "wfTraceKernel(shadow)": 32 + 16,                         # per shadow ray
        "wfShadeShadowKernel": 2 * state + 2 * (32 + 16) + 32,     # per scatter event with guide
        "histReduceKernel": 0,
    }


ISSUE_PER_SEGMENT = {"volume_mis": 1100, "volume": 400, "volume_mis_pol": 1700, "scene": 450, "scene_pol": 650}


def parseArgs():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c2pencil", "c3", "c5"])
    ap.add_argument("--batch", type=int, default=1 << 22, help="paths per step and GPU")
    ap.add_argument("--cpu-paths", type=int, default=0, help="paths of the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    return ap.parse_args()


def buildWorkload(name: str, batch: int):
    """The synthetic configurations of SURVEY.md 8d, histogram response (throughput form)."""
    import harness as H
    if name == "c1":
        cfg = H.volumeScenario(batch, resp="hist")
        desc = "C1 VolumeForwardTracer: isotropic point source in homogeneous water, sphere target, MIS, nScattering=10"
        kind = "volume_mis"
    elif name in ("c2", "c2pencil"):
        cfg = H.sceneC2(batch, resp="hist", lightKind="cone" if name == "c2" else "pencil")
        desc = ("C2 SceneForwardTracer polarized: " + ("cone" if name == "c2" else "pencil") +
                " light into sphere.stl (glass, TR) + cube.ply (detector), maxPathLength=10")
        kind = "scene_pol"
    elif name == "c3":
        cfg = H.sceneC3(batch, resp="hist")
        desc = "C3 SceneForwardTracer: suzanne.stl detector + torus.stl glass, scattering medium, sphere target guide"
        kind = "scene"
    else:
        cfg = H.sceneC5(batch, resp="hist")
        desc = "C5 SceneForwardTracer: 5000 sphere.stl modules (10x10x50 lattice) in sea water, cascade light"
        kind = "scene"
    return cfg, desc, kind


class ClockSampler:
    """Samples SM clocks and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int) -> None:
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nv = None

    def _loop(self) -> None:
        nv = self._nv
        names = {
            "hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
            "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
            "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
            "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap,
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.025)  # sampled under load; a query costs well under a millisecond

    def start(self) -> None:
        if self._nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._thread is not None:
            self._thread.join()
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def cpuBaseline(workload: str, paths: int, steps: int = 1, warmup: int = 0):
    """
    Times the CPU restatement of the reference's shader math (oracle/, the one place bench.py
    may execute it) on all host cores for a bounded sample of the same workload.
    Returns (segments/s, hits/s, cores, seconds per step, segments per step).
    """
    import oracle
    from theia_b200 import _device
    import harness as H
    # explicit thread count: torchrun exports OMP_NUM_THREADS=1
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    with _device.useDevice("cpu"):
        cfg, _, _ = buildWorkload(workload, paths)
        tr = cfg["tracer"]
        osc = oracle.OracleScene(tr.scene) if hasattr(tr, "scene") else None
        for _ in range(warmup):
            H.runOracle(cfg, 0, oracleScene=osc, threads=cores)
        segs = hits = 0
        t0 = time.perf_counter()
        for _ in range(steps):
            c = H.runOracle(cfg, 0, oracleScene=osc, threads=cores)
            segs += int(c[0])
            hits += int(c[1])
        dt = time.perf_counter() - t0
    return segs / dt, hits / dt, cores, dt / steps, segs / steps


def runReference(args) -> None:
    """--impl reference: the reference's algorithm on the host cores (CPU oracle; the GLSL /
    Vulkan reference itself cannot run without a graphics stack: "unavailable on this box")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    paths = args.cpu_paths or (1 << 19)
    segps, hitps, cores, secPerStep, segsPerStep = cpuBaseline(args.workload, paths, steps=args.steps, warmup=args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": segps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secPerStep * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "paths_per_step": paths, "l2": "n/a (CPU)"},
        "hits_per_s": hitps,
        "cpu_baseline": {"value": segps, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} steps x {paths} paths of workload {args.workload}, OpenMP over paths"},
        "e2e": {"value": segps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def runB200(args) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist
    from ctypes import byref
    from theia_b200 import _device, _lib
    from theia_b200 import pipeline as pl

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: theia_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _lib.load()

    with _device.useDevice("cuda"):
        cfg, desc, kind = buildWorkload(args.workload, args.batch)
    tr, resp, rng = cfg["tracer"], cfg["response"], cfg["rng"]
    counters = tr.enableWorkCounters()
    stages = tr.collectStages()
    pipe = pl.Pipeline(stages)
    nRNG = tr.nRNGSamples
    rng.autoAdvance = 0  # offsets are set explicitly: batch b uses offset b * nRNGSamples

    def setBatch(b: int) -> None:
        rng.setParam("offset", b * nRNG)

    # L2 flush buffer (B200 L2 = 126 MB): written between timed steps, outside the events
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    acc = torch.zeros(resp.nBins, dtype=torch.float64, device="cuda")
    reducer = None

    def deviceStep(b: int) -> None:
        """tracer + reduce kernels only, nothing leaves the device"""
        nonlocal reducer
        setBatch(b)
        pipe.update(0)
        L, keep = tr._launch(0)
        _lib.check(lib.theia_trace_forward(byref(L)))
        reducer = resp._reducer
        p = reducer._block("Params", 0)
        _lib.check(lib.theia_hist_reduce(reducer.histIn.data_ptr(), p.nHist, resp.nBins, p.normalization,
                                         reducer.histOut.data_ptr(), 1, _device.currentStream("cuda")))
        acc.add_(reducer.histOut)  # also warms the accumulate kernel (lazy module load)

    # ---------------- warm-up
    step = 0
    tWarm = time.perf_counter()
    nWarm = 0
    # W warm-up steps, then keep going until the GPU has been busy for a second (clock ramp)
    while nWarm < max(args.warmup, 3) or time.perf_counter() - tWarm < 1.0:
        deviceStep(step * world + rank)
        step += 1
        nWarm += 1
        if nWarm % 4 == 0:
            torch.cuda.synchronize()
    if world > 1:
        dist.all_reduce(acc)  # untimed: brings the NCCL communicator up before the timed region
    torch.cuda.synchronize()
    counters.zero_()
    acc.zero_()
    launch0, _k = tr._launch(0)
    grid = np.zeros(1, np.uint32)
    import ctypes
    g = ctypes.c_uint32()
    _lib.check(lib.theia_trace_grid_size(byref(launch0), byref(g)))

    # ---------------- timed region: K steps, device resident
    sampler = ClockSampler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    launches0 = int(lib.theia_kernel_launches())
    evs = []
    kernelEvs = []
    for _ in range(args.steps):
        # host side of the step (parameter update, launch description) happens before the
        # timed region: "value" is the device time with everything resident
        setBatch(step * world + rank)
        pipe.update(0)
        L, keep = tr._launch(0)
        p = reducer._block("Params", 0)
        if not os.environ.get("THEIA_BENCH_NOFLUSH"):  # diagnostic switch only
            flush.fill_(step & 0xFF)  # untimed L2 flush
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        _lib.check(lib.theia_trace_forward(byref(L)))
        e1.record()
        _lib.check(lib.theia_hist_reduce(reducer.histIn.data_ptr(), p.nHist, resp.nBins, p.normalization,
                                         reducer.histOut.data_ptr(), 1, _device.currentStream("cuda")))
        acc.add_(reducer.histOut)
        e2.record()
        evs.append((e0, e2))
        kernelEvs.append((e0, e1))
        step += 1
    launches = int(lib.theia_kernel_launches()) - launches0
    eA, eB = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    eA.record()
    if world > 1:
        dist.all_reduce(acc)  # the one collective of the run: the light curve
    eB.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    lightSum = float(acc.sum().item())
    ms = sum(a.elapsed_time(b) for a, b in evs) + eA.elapsed_time(eB)
    kernelMs = sum(a.elapsed_time(b) for a, b in kernelEvs) / args.steps
    work = counters.cpu().numpy().astype(np.int64)
    segs, hits, shadow, paths = (int(x) for x in work)

    # ---------------- end to end through the public API, host buffers
    counters.zero_()
    torch.cuda.synchronize()
    result = None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        setBatch(step * world + rank)
        pipe.run(0)  # update, launch assembly, tracer, reduce, D2H of the light curve, sync
        result = resp.result(0)
        step += 1
    e2eSec = time.perf_counter() - t0
    e2eSegs = int(counters.cpu().numpy()[0])
    h2d = ctypes.sizeof(launch0) + 512  # launch description + parameter blocks -> kernel argument space

    # ---------------- per kernel durations, live (CUDA events around every launch inside the library)
    counters.zero_()
    torch.cuda.synchronize()
    _lib.check(lib.theia_profile_enable(1))
    for _ in range(args.steps):
        if not os.environ.get("THEIA_BENCH_NOFLUSH"):
            flush.fill_(step & 0xFF)
        deviceStep(step * world + rank)
        step += 1
    torch.cuda.synchronize()
    nCls = 8
    clsMs = (ctypes.c_double * nCls)()
    clsN = (ctypes.c_uint64 * nCls)()
    lib.theia_profile_read(clsMs, clsN, nCls)
    _lib.check(lib.theia_profile_enable(0))
    profWork = counters.cpu().numpy().astype(np.int64)
    kernels = {}
    for k in range(nCls):
        if clsN[k]:
            kernels[lib.theia_profile_class_name(k).decode()] = {"launches": int(clsN[k]), "ms": float(clsMs[k])}
    d2h = resp.nBins * 4

    stats = torch.tensor([ms, float(segs), float(hits), float(shadow), float(paths), e2eSec, float(e2eSegs)],
                         dtype=torch.float64, device="cuda")
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms, e2eSec = float(mx[0]), float(mx[5])
        segs, hits, shadow, paths, e2eSegs = (int(sm[i]) for i in (1, 2, 3, 4, 6))
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # which ray query kernel the library picked for this scene (wavefront scene tracer only)
    traceChoice = None
    if hasattr(tr, "scene") and getattr(tr, "wavefront", False):
        cp, cs = ctypes.c_int(-1), ctypes.c_int(-1)
        lib.theia_get_trace_choice(tr.scene.tlas.data_ptr(), args.batch, int(tr.targetGuide is not None), byref(cp), byref(cs))
        names = {-1: "undecided", 0: "phased persistent", 1: "one thread per ray"}
        traceChoice = {"primary": names.get(cp.value), "shadow": names.get(cs.value) if tr.targetGuide is not None else None}

    # ---------------- report (rank 0)
    sec = ms * 1e-3
    value = segs / sec
    polarized = tr.polarized
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    hbmPeak = float(peaks.get("hbm_gbs", 6650.0))
    peakSrc = "measured (MEASURED_PEAKS.json)" if peaks else "fallback"
    segsPerStep = segs / (args.steps * world)
    # units each kernel class processed during the profiled pass (rank 0)
    pSegs, _pHits, pShadow, pPaths = (int(x) for x in profWork)
    units = {"volumeForwardKernel": pSegs, "sceneForwardKernel": pSegs, "wfGenerateKernel": pPaths,
             "wfTraceKernel": pSegs, "wfShadeKernel": pSegs, "wfTraceKernel(shadow)": pShadow,
             "wfShadeShadowKernel": pShadow // 2, "histReduceKernel": 0}
    perUnit = algorithmicBytes(polarized)
    totalMs = sum(k["ms"] for k in kernels.values()) or 1.0
    for name, k in kernels.items():
        k["share"] = k["ms"] / totalMs
        k["ms_per_launch"] = k["ms"] / k["launches"]
        k["algorithmic_bytes_per_launch"] = perUnit[name] * units[name] / k["launches"]
        k["achieved_gbs"] = perUnit[name] * units[name] / (k["ms"] * 1e-3) / 1e9
    dominant = max(kernels, key=lambda n: kernels[n]["ms"])
    dom = kernels[dominant]
    achievedGBs = dom["achieved_gbs"]
    traffic, ncuStatic = None, None
    try:  # figures of the dominant kernel from the committed ncu --set full capture (not measured live)
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            ncuStatic = json.load(f).get(args.workload, {}).get(dominant)
        traffic = ncuStatic.get("dram_bytes") if ncuStatic else None
    except Exception:
        pass
    smMHz = clocks.get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
    sms = _lib.smCount()
    issuePeak = sms * 128 * smMHz * 1e6
    issuePerSeg = ISSUE_PER_SEGMENT[kind + ("_pol" if polarized and kind == "volume_mis" else "")]
    issueAchieved = issuePerSeg * segsPerStep / (kernelMs * 1e-3)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": nWarm,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "description": desc, "paths_per_step_per_gpu": args.batch,
                   "segments_per_path": segs / max(paths, 1), "grid_ctas": int(g.value), "block": 128,
                   "rng": "Philox4x32-10, disjoint counter ranges per batch", "l2": "flushed between timed steps (256 MiB write)",
                   "parallelism": f"dp{world}: batches dealt round robin, one NCCL all-reduce of the light curve",
                   "ray_query_kernel": traceChoice},
        "hits_per_s": hits / sec, "shadow_rays_per_s": shadow / sec, "paths_per_s": paths / sec,
        "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achievedGBs, "peak": hbmPeak, "unit": "GB/s",
                     "frac": achievedGBs / hbmPeak, "traffic": traffic, "peak_source": peakSrc,
                     "ms_per_launch": dom["ms_per_launch"], "share_of_step": dom["share"],
                     "algorithmic_bytes_per_launch": dom["algorithmic_bytes_per_launch"],
                     "tracer_ms_per_step": kernelMs, "kernels": kernels, "ncu_capture": ncuStatic,
                     "note": "dominant kernel by live CUDA-event time; it is latency/issue bound, not HBM bound "
                             "(see roofline_issue, profiles/ and DESIGN.md section 6)"},
        "roofline_issue": {"bound": "sm_issue", "achieved": issueAchieved, "peak": issuePeak, "unit": "lane-instr/s",
                           "frac": issueAchieved / issuePeak, "declared_instr_per_segment": issuePerSeg,
                           "sm_mhz": smMHz, "sms": sms,
                           "note": "declared physics instructions only; traversal instructions come on top (ncu)"},
        "clocks": clocks,
        "e2e": {"value": e2eSegs / e2eSec, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2eSec * 1e3 / args.steps},
        "gpu_launches": launches,
        "light_curve_sum": lightSum,
    }
    if not args.no_cpu and world == 1:
        try:
            paths_cpu = args.cpu_paths or (1 << 20)
            segps, hitps, cores, secStep, _ = cpuBaseline(args.workload, paths_cpu, steps=1, warmup=0)
            # grow the sample until it is about 10 s of CPU work
            if secStep < 5.0:
                n = int(min(max(10.0 / max(secStep, 1e-3), 1.0), 16))
                segps, hitps, cores, secStep, _ = cpuBaseline(args.workload, paths_cpu, steps=n, warmup=0)
                paths_cpu *= n
            line["cpu_baseline"] = {"value": segps, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{paths_cpu} paths of workload {args.workload} (CPU oracle, OpenMP over paths)",
                                    "hits_per_s": hitps}
        except Exception as err:  # the baseline must not lose the GPU number
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"failed: {err}"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main() -> None:
    args = parseArgs()
    if args.impl == "reference":
        runReference(args)
    else:
        runB200(args)


if __name__ == "__main__":
    main()
