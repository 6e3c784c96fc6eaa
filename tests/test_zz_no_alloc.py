"""
theia_tlas_query_static: the ray query seam without any allocation (scratch memory in a static device array of
the library, no counterpart in the reference). Its output must
equal theia_tlas_query's, bit for bit (ids, t, u, v: the parity bar of every ray query test here), on two calls in
a row that share the static array - the second with other rays and a smaller count, so stale contents of the
first call would show.

Collected last on purpose (file name).
"""

import numpy as np
import pytest
import torch

import harness as H
from theia_b200 import _device, _lib

MAX_RAYS = 1 << 18


def _rays(seed, n):
    rng = np.random.default_rng(seed)
    rays = np.zeros((n, 8), np.float32)
    rays[:, :3] = rng.uniform(-2.0, 8.0, (n, 3))
    d = rng.normal(size=(n, 3))
    rays[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays[:, 3] = 1000.0
    rays[:100, 4:7] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, 100)]  # axis aligned
    return rays


@pytest.mark.gpu
def test_static_scratch_ray_query_equals_the_allocating_one(cuda):
    lib = _lib.load()
    with _device.useDevice("cuda"):
        cfg = H.sceneC2(256)
    tlas = cfg["scene"].tlas.data_ptr()
    for seed, n in ((11, 200000), (12, 70001)):
        rd = torch.from_numpy(_rays(seed, n)).cuda()
        ids = [torch.full((n, 4), -7, dtype=torch.int32, device="cuda") for _ in range(2)]
        tuv = [torch.full((n, 4), -7.0, dtype=torch.float32, device="cuda") for _ in range(2)]
        torch.cuda.synchronize()
        _lib.check(lib.theia_tlas_query(tlas, rd.data_ptr(), n, ids[0].data_ptr(), tuv[0].data_ptr(), None, 0))
        _lib.check(lib.theia_tlas_query_static(tlas, rd.data_ptr(), n, ids[1].data_ptr(), tuv[1].data_ptr(), 0))
        torch.cuda.synchronize()
        a, b = ids[0].cpu().numpy(), ids[1].cpu().numpy()
        assert np.count_nonzero(a[:, 0] >= 0) > n // 100
        assert np.array_equal(a, b), f"{np.count_nonzero((a != b).any(1))} rays differ"
        assert np.array_equal(tuv[0].cpu().numpy().view(np.uint32), tuv[1].cpu().numpy().view(np.uint32))
    # beyond the static array: rejected with a clear error, nothing launched
    rc = lib.theia_tlas_query_static(tlas, rd.data_ptr(), MAX_RAYS + 1, ids[0].data_ptr(), tuv[0].data_ptr(), 0)
    assert rc != 0 and b"at most" in lib.theia_last_error()
