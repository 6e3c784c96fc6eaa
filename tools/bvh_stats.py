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
