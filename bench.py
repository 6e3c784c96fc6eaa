"""
bench.py - photon path segments/s of the forward tracing hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3] [--batch P] [--secondary c1,c2,c5]
    python bench.py --impl reference ...      # CPU arm: the oracle on the host cores

A step = one batch of P paths through the tracer kernels + the histogram reduce kernel. The headline workload is C3
(the largest single GPU configuration of BASELINE.json: meshed scene, scattering medium, target guide); C1, C2 and C5
are measured in the same run and reported under "secondary" (C5 is the detector array the 1e10 segments/s target of
BASELINE.json is stated on).
N > 1: launched by torchrun, one rank per GPU; every rank traces its own batches (batch b = step * world + rank, its
Philox offset b * 4 * nRNGSamples: disjoint counter ranges), no data-path collective, and ONE NCCL all-reduce of the
accumulated light curve at the end of the run (inside the timed region).

Timing: CUDA events on the launching stream around each step, L2 flushed (untimed) between steps, max over ranks.
"value" has all inputs resident in HBM (scene, media; parameters travel as kernel arguments); "e2e" goes through the
public pipeline API (stage updates, launch assembly, tracer, reduce, device->host copy of the light curve into pinned
memory, synchronise) by wall clock.

roofline: SURVEY.md 8d. For every kernel class three algorithmic figures are formed from declared per-unit constants
 and counts measured in this run - lane instructions (SM issue), bytes fetched through L2 (BVH
nodes / triangles / instances, table samples) and HBM bytes (path state, rays, results) - each divided by the kernel's
live CUDA-event time and by the matching peak (issue: SMs x 128 lanes x the SM clock sampled during the run; L2: read
bandwidth measured in this run by theia_l2_probe; HBM: MEASURED_PEAKS.json). "achieved" is the largest of the three
fractions of the dominant kernel, "bound" says which. The visit counts of the ray queries come from an instrumented
step of the same workload (theia_trace_stats_enable), not from a constant.
"""

from __future__ import annotations

import argparse
import ctypes
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "photon path segments/s"
UNIT = "segments/s"
RNG_STRIDE = 4  # draw i of a stream reads Philox block offset + 4 i: batches are 4 * nRNGSamples blocks apart

# ---- declared algorithmic constants
PHILOX_DRAW = 72                                   # lane instructions per draw
NODE_INSTR, TRI_INSTR, INST_INSTR = 130, 45, 30    # per node visited / triangle tested / instance entered
NODE_BYTES, TRI_BYTES, INST_BYTES = 80, 48, 64     # bytes fetched through L2 per visit (bvh.cuh record sizes)
HIT_BYTES, TABLE_BYTES = 160, 96                   # per committed hit (processRayQuery), table samples per segment
SEGMENT_INSTR = {"volume_mis": 1100, "volume": 400, "volume_mis_pol": 1700, "scene": 450, "scene_pol": 650}
GUIDE_EVENT_INSTR = 1200                           # MIS scatter event with a target guide (two shadow rays)
GENERATE_INSTR = {"c1": 0, "c2": 4 * PHILOX_DRAW + 250, "c2pencil": 2 * PHILOX_DRAW + 150, "c3": 3 * PHILOX_DRAW + 200,
                  "c5": 9 * PHILOX_DRAW + 600}    # sampleWavelength + sampleLight + createRay (cascade: ~3 gamma trials)


def stateBytes(polarized: bool) -> int:
    """dense path state per entry: 19 words (26 polarised, incl. the path number)"""
    return 4 * (26 if polarized else 19)


def parseArgs():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c1", "c2", "c2pencil", "c3", "c5"])
    ap.add_argument("--batch", type=int, default=1 << 22, help="paths per step and GPU")
    ap.add_argument("--secondary", default="c1,c2,c5", help="workloads measured next to the headline one ('' = none)")
    ap.add_argument("--secondary-steps", type=int, default=10)
    ap.add_argument("--e2e-small-batch", type=int, default=1 << 20, help="also report e2e at this batch size (0 = off)")
    ap.add_argument("--cpu-paths", type=int, default=0, help="paths per step of the CPU sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    return ap.parse_args()


DESCRIPTIONS = {
    "c1": "C1 VolumeForwardTracer: isotropic point source in homogeneous water, sphere target, MIS, nScattering=10",
    "c2": "C2 SceneForwardTracer polarized: cone light into sphere.stl (glass, TR) + cube.ply (detector), maxPathLength=10",
    "c2pencil": "C2 SceneForwardTracer polarized: pencil light into sphere.stl (glass, TR) + cube.ply (detector), maxPathLength=10",
    "c3": "C3 SceneForwardTracer: suzanne.stl detector + torus.stl glass, scattering medium, sphere target guide, maxPathLength=10",
    "c5": "C5 SceneForwardTracer: 5000 sphere.stl modules (10x10x50 lattice) in sea water, cascade light, maxPathLength=10",
}


def buildWorkload(name: str, batch: int):
    """The synthetic configurations of SURVEY.md 8d, histogram response (throughput form)."""
    import harness as H
    if name == "c1":
        return H.volumeScenario(batch, resp="hist"), "volume_mis"
    if name in ("c2", "c2pencil"):
        return H.sceneC2(batch, resp="hist", lightKind="cone" if name == "c2" else "pencil"), "scene_pol"
    if name == "c3":
        return H.sceneC3(batch, resp="hist"), "scene"
    return H.sceneC5(batch, resp="hist"), "scene"


def staticConfig(args, world: int, name: str) -> dict:
    """What both arms (B200, reference) measure: the same dict on either line."""
    return {"workload": name, "description": DESCRIPTIONS[name], "paths_per_step_per_gpu": args.batch, "block": 128,
            "rng": f"Philox4x32-10, key per path, batch b at offset b * {RNG_STRIDE} * nRNGSamples (disjoint counter ranges)",
            "l2": "flushed between timed steps (256 MiB write)",
            "parallelism": f"dp{world}: batches dealt round robin, one NCCL all-reduce of the light curve"}


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
    Times the CPU restatement of the reference's shader math (oracle/, the one place bench.py may execute it) on all
    host cores for a bounded sample of the same workload.
    Returns (segments/s, hits/s, cores, seconds per step, segments per step).
    """
    import oracle
    from theia_b200 import _device
    import harness as H
    # explicit thread count: torchrun exports OMP_NUM_THREADS=1
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    with _device.useDevice("cpu"):
        cfg, _ = buildWorkload(workload, paths)
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
    return segs / dt, hits / dt, cores, dt / max(steps, 1), segs / max(steps, 1)


def cpuSamplePaths(workload: str, nSteps: int, budgetSeconds: float, limit: int) -> int:
    """Paths per step so that nSteps steps of the CPU arm take about budgetSeconds (from a warmed up pilot of 2^17 paths)."""
    pilot = 1 << 17
    _, _, _, sec, _ = cpuBaseline(workload, pilot, steps=1, warmup=1)
    n = int(pilot * budgetSeconds / max(nSteps, 1) / max(sec, 1e-4))
    p = 1 << 14
    while p * 2 <= min(n, limit):
        p *= 2
    return p


def runReference(args) -> None:
    """--impl reference: the reference's algorithm on the host cores (CPU oracle; the GLSL / Vulkan reference itself
    cannot run without a graphics stack: "unavailable on this box"). Same workload and config as the B200 arm; a step
    is a bounded sample of it (the metric is a rate), sized so that the whole run takes a few minutes."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    nSteps = args.steps + args.warmup
    paths = args.cpu_paths or cpuSamplePaths(args.workload, nSteps, 150.0, args.batch)
    segps, hitps, cores, secPerStep, segsPerStep = cpuBaseline(args.workload, paths, steps=args.steps, warmup=args.warmup)
    sample = (f"{args.steps} steps x {paths} paths of workload {args.workload} (of {args.batch} per step on the B200 arm), "
              f"CPU oracle, OpenMP over paths on {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": segps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secPerStep * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": staticConfig(args, world, args.workload),
        "hits_per_s": hitps, "sample_paths_per_step": paths,
        "cpu_baseline": {"value": segps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": segps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def l2PeakGBs(lib) -> float:
    """L2 read bandwidth of this device, measured: 32 MiB buffer read 64 times with 128 bit loads that bypass L1."""
    import torch
    from theia_b200 import _device, _lib
    buf = torch.zeros(32 << 20, dtype=torch.uint8, device="cuda")
    sink = torch.zeros(4, dtype=torch.int32, device="cuda")
    best = 0.0
    for it in range(9):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.theia_l2_probe(buf.data_ptr(), buf.numel(), 64, sink.data_ptr(), _device.currentStream("cuda")))
        e1.record()
        torch.cuda.synchronize()
        if it:  # the first pass brings the buffer into L2 and loads the module
            best = max(best, buf.numel() * 64 / (e0.elapsed_time(e1) * 1e-3) / 1e9)
    return best


def measure(name: str, args, steps: int, lib, *, detail: bool, batch: int | None = None) -> dict:
    """
    Warm-up, `steps` device timed steps, `steps` end to end steps of one workload on this rank; with detail=True also
    a pass with per kernel events and an instrumented step (visit counts). Collective calls (N > 1) are made by every
    rank in the same order. Returns the raw figures; rank 0 turns them into the report.
    """
    import numpy as np
    import torch
    import torch.distributed as dist
    from ctypes import byref
    from theia_b200 import _device, _lib
    from theia_b200 import pipeline as pl

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    batch = batch or args.batch
    with _device.useDevice("cuda"):
        cfg, kind = buildWorkload(name, batch)
    tr, resp, rng = cfg["tracer"], cfg["response"], cfg["rng"]
    counters = tr.enableWorkCounters()
    pipe = pl.Pipeline(tr.collectStages())
    nRNG = tr.nRNGSamples
    rng.autoAdvance = 0  # offsets are set explicitly

    def setBatch(b: int) -> None:
        rng.setParam("offset", b * RNG_STRIDE * nRNG)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # L2 flush buffer (L2 = 126 MB)
    acc = torch.zeros(resp.nBins, dtype=torch.float64, device="cuda")
    stream = _device.currentStream("cuda")

    def deviceStep(b: int) -> None:
        """tracer + reduce kernels only, nothing leaves the device"""
        setBatch(b)
        pipe.update(0)
        L, keep = tr._launch(0)
        _lib.check(lib.theia_trace_forward(byref(L)))
        red = resp._reducer
        p = red._block("Params", 0)
        _lib.check(lib.theia_hist_reduce(red.histIn.data_ptr(), p.nHist, resp.nBins, p.normalization,
                                         red.histOut.data_ptr(), 1, stream))
        acc.add_(red.histOut)

    wavefront = hasattr(tr, "scene") and getattr(tr, "wavefront", False)
    guide = wavefront and tr.targetGuide is not None

    def traceChoice():
        if not wavefront:
            return None
        cp, cs = ctypes.c_int(-1), ctypes.c_int(-1)
        lib.theia_get_trace_choice(tr.scene.tlas.data_ptr(), batch, int(guide), byref(cp), byref(cs))
        return cp.value, cs.value

    # ---------------- warm-up: W steps (at least 3), at least a second of GPU work (clock ramp), and until the
    # library has picked its ray query kernels (nine batches, api.cu)
    step, nWarm, tWarm = 0, 0, time.perf_counter()
    while True:
        undecided = wavefront and nWarm < 16 and min(traceChoice()[0], traceChoice()[1] if guide else 0) < 0
        if nWarm >= max(args.warmup, 3) and time.perf_counter() - tWarm >= 1.0 and not undecided:
            break
        deviceStep(step * world + rank)
        step += 1
        nWarm += 1
        if nWarm % 3 == 0:
            torch.cuda.synchronize()
    if world > 1:
        dist.all_reduce(acc)  # untimed: brings the NCCL communicator up before the timed region
    torch.cuda.synchronize()
    counters.zero_()
    acc.zero_()
    launch0, _k = tr._launch(0)
    g = ctypes.c_uint32()
    _lib.check(lib.theia_trace_grid_size(byref(launch0), byref(g)))

    # ---------------- timed region: K steps, device resident
    sampler = ClockSampler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    launches0 = int(lib.theia_kernel_launches())
    evs, kernelEvs = [], []
    red = resp._reducer
    for _ in range(steps):
        # host side of the step (parameter update, launch description) happens before the timed region: "value" is
        # the device time with everything resident
        setBatch(step * world + rank)
        pipe.update(0)
        L, keep = tr._launch(0)
        p = red._block("Params", 0)
        flush.fill_(step & 0xFF)  # untimed L2 flush
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        _lib.check(lib.theia_trace_forward(byref(L)))
        e1.record()
        _lib.check(lib.theia_hist_reduce(red.histIn.data_ptr(), p.nHist, resp.nBins, p.normalization,
                                         red.histOut.data_ptr(), 1, stream))
        acc.add_(red.histOut)
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
    out = {"name": name, "kind": kind, "polarized": bool(tr.polarized), "batch": batch, "steps": steps, "warmup_run": nWarm,
           "grid": int(g.value), "clocks": clocks, "launches": launches, "guide": bool(guide),
           "light_curve_sum": float(acc.sum().item()),
           "ms": sum(a.elapsed_time(b) for a, b in evs) + eA.elapsed_time(eB),
           "tracer_ms_per_step": sum(a.elapsed_time(b) for a, b in kernelEvs) / steps,
           "trace_choice": traceChoice()}
    out["segs"], out["hits"], out["shadow"], out["paths"] = (int(x) for x in counters.cpu().numpy().astype(np.int64))

    # ---------------- end to end through the public API, host buffers: ShardedRun (theia_b200/distributed.py), the call a
    # user of N GPUs makes: ONE run of steps x N batches. Per batch on its rank: parameter update and launch assembly on
    # the host, tracer, reduce kernel, device side accumulation, D2H of the batch's light curve handed to processFn
    # (synchronises); at the end one all-reduce of the accumulated curve (N > 1) and its D2H.
    from theia_b200.distributed import ShardedRun, batchOffset

    def e2eRun(tracer, nRng, n) -> tuple[float, int]:
        sharded = ShardedRun(tracer, offset0=batchOffset(0, step * world, nRng, RNG_STRIDE))  # rank r: batches = r mod N
        sharded.run(2 * world, processFn=lambda b, curve: None)  # untimed: the API path's own first-call costs
        sharded.offset0 = batchOffset(0, (step + 2) * world, nRng, RNG_STRIDE)
        counters.zero_()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        seen = []
        t0 = time.perf_counter()
        res = sharded.run(n * world, processFn=lambda b, curve: seen.append(float(curve[0])))
        dt = time.perf_counter() - t0
        assert res.lightCurve is not None and len(seen) == n
        return dt, int(counters.cpu().numpy()[0])

    out["e2e_sec"], out["e2e_segs"] = e2eRun(tr, nRNG, steps)
    step += steps + 2
    out["h2d"] = ctypes.sizeof(launch0) + 512  # launch description + parameter blocks -> kernel argument space
    out["d2h"] = resp.nBins * 4  # the batch's light curve (+ once per run the float64 sum over all batches)

    # ---------------- per kernel durations, live (CUDA events around every launch inside the library), then one
    # instrumented step for the visit counts of the ray queries
    if detail:
        counters.zero_()
        torch.cuda.synchronize()
        _lib.check(lib.theia_profile_enable(1))
        for _ in range(steps):
            flush.fill_(step & 0xFF)
            deviceStep(step * world + rank)
            step += 1
        torch.cuda.synchronize()
        nCls = 8
        clsMs, clsN = (ctypes.c_double * nCls)(), (ctypes.c_uint64 * nCls)()
        lib.theia_profile_read(clsMs, clsN, nCls)
        _lib.check(lib.theia_profile_enable(0))
        out["prof_work"] = [int(x) for x in counters.cpu().numpy().astype(np.int64)]
        out["kernels"] = {lib.theia_profile_class_name(k).decode(): {"launches": int(clsN[k]), "ms": float(clsMs[k])}
                          for k in range(nCls) if clsN[k]}
        if wavefront:
            stats = torch.zeros(8, dtype=torch.int64, device="cuda")
            _lib.check(lib.theia_trace_stats_enable(stats.data_ptr()))
            counters.zero_()
            deviceStep(step * world + rank)
            step += 1
            torch.cuda.synchronize()
            _lib.check(lib.theia_trace_stats_enable(None))
            out["visits"] = [int(x) for x in stats.cpu().numpy()]
            # shadow rays the path tracing algorithm asks for (two per MIS event) vs those that had to be traced
            # (the others cannot reach the target: wfTargetBoxKernel), both of the instrumented step
            issued = int(counters.cpu().numpy()[2])
            out["shadow_traced_fraction"] = (out["visits"][7] / issued) if issued else 1.0

    # ---------------- the reference's own batch size (2^20 paths, examples/01) end to end, for the launch overhead
    if detail and args.e2e_small_batch and args.e2e_small_batch < batch:
        del pipe, cfg
        with _device.useDevice("cuda"):
            cfgS, _ = buildWorkload(name, args.e2e_small_batch)
        trS, respS, rngS = cfgS["tracer"], cfgS["response"], cfgS["rng"]
        countersS = trS.enableWorkCounters()
        pipeS = pl.Pipeline(trS.collectStages())
        rngS.autoAdvance = 0
        nS = trS.nRNGSamples
        for k in range(12):  # warm-up incl. the kernel choice for this batch size
            rngS.setParam("offset", (k * world + rank) * RNG_STRIDE * nS)
            pipeS.run(0)
        counters = countersS
        dt, segs = e2eRun(trS, nS, 4 * steps)
        out["e2e_small"] = {"paths_per_step_per_gpu": args.e2e_small_batch, "value": segs / dt, "ms_per_step": dt * 1e3 / (4 * steps),
                            "steps": 4 * steps}
        out["e2e_small_segs"], out["e2e_small_sec"] = segs, dt
    return out


def rooflines(m: dict, peaks: dict, l2Peak: float, sms: int) -> dict:
    """Per kernel class: algorithmic lane instructions / L2 bytes / HBM bytes of the profiled pass over its live time,
    each as a fraction of its peak. Needs measure(..., detail=True)."""
    pol, kind, name = m["polarized"], m["kind"], m["name"]
    segs, _hits, shadow, paths = m["prof_work"]
    v = m.get("visits") or [0] * 8
    perRay = lambda k, c: (v[4 * c + k] / v[4 * c + 3]) if v[4 * c + 3] else 0.0  # noqa: E731
    nodes, tris, inst = ([perRay(k, c) for c in (0, 1)] for k in (0, 1, 2))
    state = stateBytes(pol)
    physics = SEGMENT_INSTR[kind + ("_pol" if pol and kind == "volume_mis" else "")]
    events = shadow // 2
    traced = int(round(shadow * m.get("shadow_traced_fraction", 1.0)))  # shadow rays that reached the ray query kernel
    # class -> (units, lane instructions, L2 bytes, HBM bytes) per unit
    model = {
        "volumeForwardKernel": (segs, physics, TABLE_BYTES, 0),
        "sceneForwardKernel": (segs, physics + NODE_INSTR * nodes[0] + TRI_INSTR * tris[0] + INST_INSTR * inst[0],
                               NODE_BYTES * nodes[0] + TRI_BYTES * tris[0] + INST_BYTES * inst[0] + TABLE_BYTES + HIT_BYTES, 0),
        "wfGenerateKernel": (paths, GENERATE_INSTR[name], TABLE_BYTES, state + 32),
        "wfTraceKernel": (segs, NODE_INSTR * nodes[0] + TRI_INSTR * tris[0] + INST_INSTR * inst[0],
                          NODE_BYTES * nodes[0] + TRI_BYTES * tris[0] + INST_BYTES * inst[0], 32 + 16),
        # (every segment is charged a committed hit on the L2 side: an upper bound of the declared figure)
        "wfShadeKernel": (segs, physics, TABLE_BYTES + HIT_BYTES, 2 * state + 2 * 32 + 16),
        "wfTraceKernel(shadow)": (traced, NODE_INSTR * nodes[1] + TRI_INSTR * tris[1] + INST_INSTR * inst[1],
                                  NODE_BYTES * nodes[1] + TRI_BYTES * tris[1] + INST_BYTES * inst[1], 32 + 16),
        "wfShadeShadowKernel": (events, GUIDE_EVENT_INSTR, 2 * HIT_BYTES + TABLE_BYTES, 2 * state + 2 * (32 + 16) + 32),
        "histReduceKernel": (0, 0, 0, 0),
    }
    smMHz = m["clocks"].get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
    peak = {"issue": sms * 128 * smMHz * 1e6, "l2": l2Peak * 1e9, "hbm": float(peaks.get("hbm_gbs", 6650.0)) * 1e9}
    unit = {"issue": "lane-instr/s", "l2": "B/s", "hbm": "B/s"}
    kernels = {}
    totalMs = sum(k["ms"] for k in m["kernels"].values()) or 1.0
    for cls, k in m["kernels"].items():
        units, instr, l2b, hbmb = model[cls]
        sec = k["ms"] * 1e-3
        fr = {}
        for b, per in (("issue", instr), ("l2", l2b), ("hbm", hbmb)):
            ach = per * units / sec if sec > 0 else 0.0
            fr[b] = {"achieved": ach, "peak": peak[b], "unit": unit[b], "frac": ach / peak[b] if peak[b] else 0.0,
                     "algorithmic_per_unit": per}
        bound = max(fr, key=lambda b: fr[b]["frac"])
        kernels[cls] = {"launches": k["launches"], "ms": k["ms"], "ms_per_launch": k["ms"] / k["launches"],
                        "share": k["ms"] / totalMs, "units": units, "units_per_launch": units / k["launches"],
                        "bound": bound, "frac": fr[bound]["frac"], "fractions": fr,
                        "algorithmic_hbm_bytes_per_launch": hbmb * units / k["launches"]}
    dominant = max(kernels, key=lambda n: kernels[n]["ms"])
    d = kernels[dominant]
    b = d["bound"]
    traffic = None
    try:  # DRAM bytes per launch of the same launch class from the committed ncu --set full capture of this workload
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get(name, {}).get(dominant)
        if t and t.get("launch_class") == "average over one step":
            traffic = t.get("dram_bytes_per_launch")
    except Exception:
        pass
    return {"bound": b, "kernel": dominant, "achieved": d["fractions"][b]["achieved"], "peak": d["fractions"][b]["peak"],
            "unit": d["fractions"][b]["unit"], "frac": d["frac"], "traffic": traffic,
            "fractions": {x: d["fractions"][x]["frac"] for x in ("issue", "l2", "hbm")},
            "share_of_step": d["share"], "ms_per_launch": d["ms_per_launch"],
            "peaks": {"issue_lane_instr_per_s": peak["issue"], "sm_mhz": smMHz, "sms": sms,
                      "l2_read_gbs": l2Peak, "l2_source": "measured in this run (theia_l2_probe, 32 MiB x 64 passes)",
                      "hbm_gbs": peak["hbm"] / 1e9,
                      "hbm_source": "measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md)"},
            "visits_per_ray": {"primary": {"nodes": nodes[0], "triangles": tris[0], "instances": inst[0]},
                               "shadow": {"nodes": nodes[1], "triangles": tris[1], "instances": inst[1]}},
            "shadow_rays_traced_fraction": m.get("shadow_traced_fraction", 1.0),
            "kernels": kernels,
            "note": "achieved = algorithmic work (declared per-unit constants x counts of this run, visit counts from an "
                    "instrumented step) / live CUDA-event time of the dominant kernel; bound = the largest of the issue / "
                    "L2 / HBM fractions (SURVEY.md 8d)"}


def reduceOverRanks(m: dict) -> dict:
    """max of the times, sum of the work over ranks (collective; every rank calls it)"""
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1:
        return m
    keys = ["ms", "e2e_sec", "segs", "hits", "shadow", "paths", "e2e_segs", "e2e_small_segs", "e2e_small_sec"]
    t = torch.tensor([float(m.get(k, 0.0)) for k in keys], dtype=torch.float64, device="cuda")
    mx, sm = t.clone(), t.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    for i, k in enumerate(keys):
        m[k] = float(mx[i]) if k in ("ms", "e2e_sec", "e2e_small_sec") else int(sm[i])
    if "e2e_small" in m:  # whole job: segments of all ranks over the slowest rank's time
        m["e2e_small"]["value"] = m["e2e_small_segs"] / m["e2e_small_sec"]
        m["e2e_small"]["ms_per_step"] = m["e2e_small_sec"] * 1e3 / m["e2e_small"]["steps"]
    return m


KERNEL_NAMES = {-1: "undecided", 0: "phased persistent", 1: "one thread per ray", 2: "queued triangle tests"}


def summary(m: dict, args, world: int) -> dict:
    sec = m["ms"] * 1e-3
    tc = m["trace_choice"]
    return {"value": m["segs"] / sec, "unit": UNIT, "ms_per_step": m["ms"] / m["steps"], "steps": m["steps"],
            "hits_per_s": m["hits"] / sec, "shadow_rays_per_s": m["shadow"] / sec, "paths_per_s": m["paths"] / sec,
            "segments_per_path": m["segs"] / max(m["paths"], 1),
            "e2e": {"value": m["e2e_segs"] / m["e2e_sec"], "unit": UNIT, "h2d_bytes_per_step": m["h2d"],
                    "d2h_bytes_per_step": m["d2h"], "ms_per_step": m["e2e_sec"] * 1e3 / m["steps"],
                    "small_batch": m.get("e2e_small")},
            "clocks": m["clocks"], "gpu_launches": m["launches"], "grid_ctas": m["grid"],
            "ray_query_kernel": None if tc is None else {"primary": KERNEL_NAMES.get(tc[0]),
                                                         "shadow": KERNEL_NAMES.get(tc[1]) if m["guide"] else None},
            "light_curve_sum": m["light_curve_sum"]}


def runB200(args) -> None:
    import torch
    import torch.distributed as dist
    from theia_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: theia_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _lib.load()
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass

    m = reduceOverRanks(measure(args.workload, args, args.steps, lib, detail=True))
    l2Peak = l2PeakGBs(lib)
    secondary = {}
    for name in [s for s in args.secondary.split(",") if s and s != args.workload]:
        torch.cuda.empty_cache()
        s = reduceOverRanks(measure(name, args, max(3, min(args.secondary_steps, args.steps)), lib, detail=True))
        if rank == 0:
            r = rooflines(s, peaks, l2Peak, _lib.smCount())
            secondary[name] = {"description": DESCRIPTIONS[name], **summary(s, args, world),
                               "roofline": {k: r[k] for k in ("bound", "kernel", "achieved", "peak", "unit", "frac", "fractions",
                                                              "share_of_step", "ms_per_launch", "visits_per_ray")},
                               "kernel_ms_per_launch": {k: v["ms_per_launch"] for k, v in r["kernels"].items()}}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    head = summary(m, args, world)
    line = {
        "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "warmup_steps_run": m["warmup_run"], "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": staticConfig(args, world, args.workload),
        "run": {"segments_per_path": head["segments_per_path"], "grid_ctas": head["grid_ctas"],
                "ray_query_kernel": head["ray_query_kernel"], "tracer_ms_per_step": m["tracer_ms_per_step"]},
        "hits_per_s": head["hits_per_s"], "shadow_rays_per_s": head["shadow_rays_per_s"], "paths_per_s": head["paths_per_s"],
        "roofline": rooflines(m, peaks, l2Peak, _lib.smCount()),
        "clocks": head["clocks"], "e2e": head["e2e"], "gpu_launches": head["gpu_launches"],
        "light_curve_sum": head["light_curve_sum"],
        "secondary": secondary,
    }
    if not args.no_cpu and world == 1:
        try:
            paths_cpu = args.cpu_paths or cpuSamplePaths(args.workload, 1, 20.0, 8 * args.batch)
            segps, hitps, cores, secStep, _ = cpuBaseline(args.workload, paths_cpu, steps=1, warmup=0)
            line["cpu_baseline"] = {"value": segps, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{paths_cpu} paths of workload {args.workload} ({secStep:.1f} s; CPU oracle, "
                                              f"OpenMP over paths)", "hits_per_s": hitps}
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
