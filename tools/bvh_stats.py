"""Traversal statistics of the GPU BVH on the rays a workload really traces (needs a GPU).

    python tools/bvh_stats.py c2 c3 c5

Rays are rebuilt from the event tracks of the tracer (TrackRecordCallback): consecutive
vertices of a path give origin, direction and length of the ray query of that segment.
Prints nodes / triangles / instances visited per ray (theia_tlas_query statistics).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402
from theia_b200 import _device, _lib  # noqa: E402

BUILDERS = {"c2": (H.sceneC2, dict(lightKind="cone")), "c3": (H.sceneC3, dict()), "c5": (H.sceneC5, dict())}


def workloadRays(name: str, N: int = 1 << 16) -> tuple[np.ndarray, object]:
    builder, kw = BUILDERS[name]
    with _device.useDevice("cuda"):
        cfg = builder(N, resp="hist", callback="track", **kw)
    H.runGpu(cfg)
    torch.cuda.synchronize()
    tracks, lengths, codes = cfg["callback"].result(0)
    rays = []
    L = tracks.shape[1]
    for k in range(L - 1):
        ok = lengths > k + 1
        a, b = tracks[ok, k, :3], tracks[ok, k + 1, :3]
        d = b - a
        dist = np.linalg.norm(d, axis=1)
        good = dist > 1e-6
        r = np.zeros((int(good.sum()), 8), np.float32)
        r[:, :3] = a[good]
        r[:, 3] = dist[good] * 1.0001
        r[:, 4:7] = d[good] / dist[good, None]
        rays.append(r)
    return np.concatenate(rays), cfg


for name in sys.argv[1:] or ["c2", "c3", "c5"]:
    rays, cfg = workloadRays(name)
    n = len(rays)
    rd = torch.from_numpy(rays).cuda()
    ids = torch.zeros(n, 4, dtype=torch.int32, device="cuda")
    tuv = torch.zeros(n, 4, dtype=torch.float32, device="cuda")
    st = torch.zeros(4, dtype=torch.int64, device="cuda")
    _lib.check(_lib.load().theia_tlas_query(cfg["scene"].tlas.data_ptr(), rd.data_ptr(), n, ids.data_ptr(), tuv.data_ptr(),
                                            st.data_ptr(), 0))
    torch.cuda.synchronize()
    s = st.cpu().numpy()
    hit = int((ids[:, 0] >= 0).sum().item())
    print(f"{name}: {n} rays, {hit / n:.2f} hit; per ray: nodes {s[0] / n:.2f} triangles {s[1] / n:.2f} instances {s[2] / n:.2f}",
          flush=True)


def timeQuery(tlas, rays: np.ndarray, reps: int = 5) -> float:
    """kernel time (ms) of the ray query through the library's own per kernel timers"""
    import ctypes
    n = len(rays)
    rd = torch.from_numpy(np.ascontiguousarray(rays)).cuda()
    ids = torch.zeros(n, 4, dtype=torch.int32, device="cuda")
    tuv = torch.zeros(n, 4, dtype=torch.float32, device="cuda")
    lib = _lib.load()
    best = 1e9
    for _ in range(reps):
        lib.theia_profile_enable(1)
        _lib.check(lib.theia_tlas_query(tlas, rd.data_ptr(), n, ids.data_ptr(), tuv.data_ptr(), 0, 0))
        torch.cuda.synchronize()
        ms = (ctypes.c_double * 8)()
        cnt = (ctypes.c_uint64 * 8)()
        lib.theia_profile_read(ms, cnt, 8)
        lib.theia_profile_enable(0)
        best = min(best, ms[1] + ms[3])  # sceneForward class (query kernel) or wfTrace class
    return best


def mortonKey(p: np.ndarray, lo: np.ndarray, hi: np.ndarray, bits: int) -> np.ndarray:
    q = np.clip(((p - lo) / np.maximum(hi - lo, 1e-9) * (1 << bits)).astype(np.int64), 0, (1 << bits) - 1)
    key = np.zeros(len(p), np.int64)
    for b in range(bits):
        for a in range(3):
            key |= ((q[:, a] >> b) & 1) << (3 * b + a)
    return key


if os.environ.get("COHERENCE_EXPERIMENT"):
    # upper bound of what coherence can buy: every warp gets 32 copies of one ray (no divergence
    # at all) against the same rays in path order
    for name in sys.argv[1:] or ["c2", "c3", "c5"]:
        rays, cfg = workloadRays(name, 1 << 18)
        rays = rays[: len(rays) // 32 * 32].copy()
        rays[:, 3] = 1000.0
        same = rays.reshape(-1, 32, 8)[:, :1, :].repeat(32, axis=1).reshape(-1, 8)
        t0, t1 = timeQuery(cfg["scene"].tlas.data_ptr(), rays), timeQuery(cfg["scene"].tlas.data_ptr(), same)
        mode = "wavefront trace kernel" if os.environ.get("THEIA_QUERY_WAVEFRONT") else "query kernel"
        print(f"{name}: {len(rays)} rays, {mode}: {t0:.3f} ms in path order, {t1:.3f} ms with 32 identical rays per warp",
              flush=True)
    sys.exit(0)

if os.environ.get("SORT_EXPERIMENT"):
    # how much does ray coherence buy? time the (one thread per ray) query kernel on the rays
    # of a workload in path order and sorted by the point where they enter the scene box
    for name in sys.argv[1:] or ["c2", "c3", "c5"]:
        rays, cfg = workloadRays(name, 1 << 18)
        sc = cfg["scene"]
        verts = np.concatenate([np.asarray(i.worldBounds()) for i in sc.instances]) if hasattr(sc.instances[0], "worldBounds") else None
        o, d, tmax = rays[:, :3].astype(np.float64), rays[:, 4:7].astype(np.float64), rays[:, 3].astype(np.float64)
        # scene box from the hit points themselves (good enough for the experiment)
        end = o + d * tmax[:, None]
        lo, hi = np.percentile(end, 1, axis=0), np.percentile(end, 99, axis=0)
        inv = 1.0 / np.where(np.abs(d) < 1e-12, 1e-12, d)
        t0, t1 = (lo - o) * inv, (hi - o) * inv
        tin = np.clip(np.minimum(t0, t1).max(axis=1), 0.0, tmax)
        entry = o + d * tin[:, None]
        for bits in (6,):
            order = np.argsort(mortonKey(entry, lo, hi, bits), kind="stable")
            for far in (False, True):
                r = rays.copy()
                if far:
                    r[:, 3] = 1000.0  # like the tracer: the query reaches far beyond the first hit
                tu, ts = timeQuery(sc.tlas.data_ptr(), r), timeQuery(sc.tlas.data_ptr(), r[order])
                mode = "wavefront trace kernel" if os.environ.get("THEIA_QUERY_WAVEFRONT") else "query kernel"
                print(f"{name}: {len(rays)} rays tmax={'far' if far else 'hit'} {mode}: {tu:.3f} ms unsorted, {ts:.3f} ms "
                      f"sorted by entry point ({3 * bits} bit morton) -> {len(rays) / tu / 1e6:.0f} / {len(rays) / ts / 1e6:.0f} Mrays/s",
                      flush=True)
