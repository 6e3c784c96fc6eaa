"""
Host logic of the multi-GPU sharding (theia_b200/distributed.py) on CPU:
two gloo ranks deal the batches round robin, run them (through the test oracle in place of
the CUDA pipeline), all-reduce the light curve and the event counters, and must reproduce
the single-process run of the same batches.
"""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import harness as H
from theia_b200 import _device, distributed

N_PATHS = 8192
N_BATCHES = 5


def freePort() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def makeRun(group=None, perObject: bool = False):
    with _device.useDevice("cpu"):
        if perObject:  # C5 style: one light curve per detector module
            cfg = H.sceneC5(N_PATHS, resp="objhist", callback="stat", lattice=(3, 3, 3), pitch=(3.0, 3.0, 3.0), nBins=40,
                            binSize=2.5)
        else:
            cfg = H.volumeScenario(N_PATHS, resp="hist", callback="stat")
    tr = cfg["tracer"]

    def launcher(i: int) -> None:  # stands in for Pipeline.run: same stages, executed by the oracle
        H.runOracle(cfg, i)

    return cfg, distributed.ShardedRun(tr, offset0=7, group=group, launcher=launcher)


def worker(rank: int, world: int, port: int, outDir: str, perObject: bool = False) -> None:
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg, run = makeRun(perObject=perObject)
        res = run.run(N_BATCHES)
        assert res.localBatches == list(range(rank, N_BATCHES, world))
        np.save(os.path.join(outDir, f"curve{rank}.npy"), res.lightCurve)
        np.save(os.path.join(outDir, f"counters{rank}.npy"), np.array(list(res.counters.values()), np.int64))
    finally:
        dist.destroy_process_group()


def test_batches_are_dealt_round_robin():
    assert list(distributed.batchesOf(0, 2, 5)) == [0, 2, 4]
    assert list(distributed.batchesOf(1, 2, 5)) == [1, 3]
    assert sorted(b for r in range(8) for b in distributed.batchesOf(r, 8, 29)) == list(range(29))
    assert list(distributed.batchesOf(3, 4, 2)) == []
    with pytest.raises(ValueError):
        distributed.batchesOf(2, 2, 4)
    assert distributed.batchOffset(5, 3, 100) == 305
    assert distributed.batchOffset((1 << 128) - 1, 1, 2) == 1


def test_two_ranks_reproduce_the_single_process_run(tmp_path, orc):
    cfg, run = makeRun()
    single = run.run(N_BATCHES)
    assert single.localBatches == list(range(N_BATCHES))
    assert single.lightCurve.sum() > 0
    port = freePort()
    mp.spawn(worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    curves = [np.load(tmp_path / f"curve{r}.npy") for r in range(2)]
    counters = [np.load(tmp_path / f"counters{r}.npy") for r in range(2)]
    assert np.array_equal(curves[0], curves[1])  # every rank holds the reduced result
    assert np.allclose(curves[0], single.lightCurve, rtol=1e-6, atol=0.0)  # float64, order of summation only
    assert np.array_equal(counters[0], counters[1])
    assert np.array_equal(counters[0], np.array(list(single.counters.values()), np.int64))
    assert single.counters["created"] == N_BATCHES * N_PATHS


def test_two_ranks_reduce_per_object_light_curves(tmp_path, orc):
    # config C5: [module][time] histograms, one all-reduce of the whole array
    cfg, run = makeRun(perObject=True)
    single = run.run(N_BATCHES)
    assert single.lightCurve.shape == (27, 40) and single.lightCurve.sum() > 0
    port = freePort()
    mp.spawn(worker, args=(2, port, str(tmp_path), True), nprocs=2, join=True)
    curves = [np.load(tmp_path / f"curve{r}.npy") for r in range(2)]
    assert np.array_equal(curves[0], curves[1])
    assert np.allclose(curves[0], single.lightCurve, rtol=1e-5, atol=1e-9 * single.lightCurve.max())
