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
    assert distributed.batchOffset(5, 3, 100) == 1205  # batches are 4 * nRNGSamples blocks apart
    assert distributed.batchOffset(5, 3, 100, rngStride=1) == 305  # the reference's autoAdvance = nRNGSamples
    assert distributed.batchOffset((1 << 128) - 1, 1, 2, 1) == 1


@pytest.mark.parametrize("nRNG", [8, 16, 42, 53, 63, 64, 66])
def test_batches_read_disjoint_philox_blocks(nRNG):
    # draw i reads block offset + 4 i (sh/random.philox.glsl:17): with the default stride no (block, element) pair
    # is read by two batches, whatever nRNGSamples is; with the reference's stride 1 batches b and b + 2 overlap
    seen = set()
    for b in range(12):
        blocks = set(distributed.blocksOfBatch(7, b, nRNG))
        assert len(blocks) == nRNG and not (blocks & seen)
        seen |= blocks
    ref = [set(distributed.blocksOfBatch(7, b, nRNG, 1)) for b in range(4)]
    if nRNG % 4 == 0:
        assert ref[0] & ref[1] or ref[0] & ref[2]  # overlapping blocks: what the default stride avoids


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


def test_process_function_sees_every_local_batch(orc):
    # the per batch hook (the reference's PipelineScheduler processFn): results of each batch on the host, in order,
    # and their mean is the reduced light curve
    cfg, run = makeRun()
    seen = []
    res = run.run(N_BATCHES, processFn=lambda b, curve: seen.append((b, np.array(curve, dtype=np.float64))))
    assert [b for b, _ in seen] == list(range(N_BATCHES))
    assert np.allclose(np.mean([c for _, c in seen], axis=0), res.lightCurve, rtol=1e-6, atol=0.0)


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


# ------------------------------------------------------------------ C4: tasks of several pipelines, gathered hits
def makeC4(group=None):
    with _device.useDevice("cpu"):
        c = H.volumeC4(N_PATHS)

    def launcherOf(name):
        def launch(i: int) -> None:
            H.runOracle({"tracer": c["tracers"][name], "response": c["response"]}, i)
        return launch

    run = distributed.ShardedRun(c["tracers"], offset0=3, group=group, launcher={n: launcherOf(n) for n in c["tracers"]})
    return c, run


def workerC4(rank: int, world: int, port: int, outDir: str) -> None:
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        c, run = makeC4()
        res = run.runTasks(c["tasks"])
        assert res.localBatches == list(range(rank, len(c["tasks"]), world))
        np.savez(os.path.join(outDir, f"c4_{rank}.npz"), **res.lightCurves)
    finally:
        dist.destroy_process_group()


def test_two_ranks_share_the_tasks_of_several_pipelines(tmp_path, orc):
    # config C4 (examples/03 of the reference): whole tasks are dealt round robin; parameters persist per pipeline
    c, run = makeC4()
    single = run.runTasks(c["tasks"])
    assert set(single.lightCurves) == set(c["tracers"]) and sum(single.taskCounts.values()) == len(c["tasks"])
    assert all(v.sum() > 0 for v in single.lightCurves.values())
    port = freePort()
    mp.spawn(workerC4, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = [np.load(tmp_path / f"c4_{r}.npz") for r in range(2)]
    for name, want in single.lightCurves.items():
        assert np.array_equal(got[0][name], got[1][name])
        assert np.allclose(got[0][name], want, rtol=1e-6, atol=0.0), name


def workerHits(rank: int, world: int, port: int, outDir: str) -> None:
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg, run = makeHitRun()
        hits = run.gatherHits(run.run(N_BATCHES, keepHits=True))
        np.savez(os.path.join(outDir, f"hits{rank}.npz"), **hits)
    finally:
        dist.destroy_process_group()


def makeHitRun():
    with _device.useDevice("cpu"):
        cfg = H.volumeScenario(N_PATHS, resp="record", polarized=True)
    return cfg, distributed.ShardedRun(cfg["tracer"], offset0=11, launcher=lambda i: H.runOracle(cfg, i))


def test_two_ranks_gather_their_hits_in_batch_order(tmp_path, orc):
    cfg, run = makeHitRun()
    single = run.gatherHits(run.run(N_BATCHES, keepHits=True))
    assert len(single["batch"]) > 0 and np.all(np.diff(single["batch"]) >= 0)
    assert set(np.unique(single["batch"])) == set(range(N_BATCHES))
    port = freePort()
    mp.spawn(workerHits, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = [np.load(tmp_path / f"hits{r}.npz") for r in range(2)]
    for f in single:
        assert np.array_equal(got[0][f], got[1][f]), f           # every rank holds all hits
        assert got[0][f].dtype == single[f].dtype
        assert np.array_equal(got[0][f], single[f]), f           # same hits, same order as one process
