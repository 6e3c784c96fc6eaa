"""
Multi-GPU sharding of tracer runs: one process per GPU, paths are independent, so
the only exchange is one all-reduce of the accumulated light curve(s) (and of the event counters) at the end
of the run, plus - on request - one all-gather of the recorded hits.

Batch (or task) ``b`` of a run uses the Philox offset ``offset0 + b * rngStride * nRNGSamples`` and the streams
``0 .. batchSize-1``; rank ``r`` of ``R`` executes the batches ``{b : b mod R = r}``. Results therefore do not
depend on the number of GPUs except for the float summation order of the final reduction, which is done in
float64.

``rngStride``: draw ``i`` of a stream reads the Philox block ``offset + 4 i`` (the ``i << 2`` of
sh/random.philox.glsl:17, reproduced in csrc/philox.cuh), so a batch that makes ``nRNGSamples`` draws covers the
counters ``[offset, offset + 4 nRNGSamples)``. The default stride 4 therefore gives every batch - and with it
every rank - its own counter range: no (block, element) pair is read twice in a run. ``rngStride = 1`` is what the
reference does (``autoAdvance = nRNGSamples``, src/theia/random.py:278-282): batch ``b + 2`` then re-reads blocks of
batch ``b``, and draws repeat outright when ``nRNGSamples`` is a multiple of 8. Use 1 only to reproduce a
reference-style run batch for batch.

The collectives run on ``torch.distributed`` (NCCL over NVLink on the GPU box; gloo in the CPU tests of the host
logic). Nothing here computes: batches are executed by the pipelines handed in, i.e. by the CUDA library, and
the light curves are accumulated on the device without a host round trip per batch.
"""

from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Iterable

import numpy as np
import torch
import torch.distributed as dist

from theia_b200 import pipeline as pl

__all__ = ["batchesOf", "batchOffset", "blocksOfBatch", "ShardResult", "ShardedRun"]

RNG_STRIDE = 4


def batchesOf(rank: int, world: int, nBatches: int) -> range:
    """Batches rank ``rank`` of ``world`` executes: dealt round robin."""
    if not 0 <= rank < world:
        raise ValueError("rank must be in [0, world)")
    return range(rank, nBatches, world)


def batchOffset(offset0: int, batch: int, nRNGSamples: int, rngStride: int = RNG_STRIDE) -> int:
    """Philox offset of a batch: 128 bit counter arithmetic wraps like the device's."""
    return (offset0 + batch * rngStride * nRNGSamples) & ((1 << 128) - 1)


def blocksOfBatch(offset0: int, batch: int, nRNGSamples: int, rngStride: int = RNG_STRIDE) -> range:
    """Philox blocks (counter values) a batch may read: draw i < nRNGSamples reads block offset + 4 i."""
    off = offset0 + batch * rngStride * nRNGSamples
    return range(off, off + 4 * nRNGSamples, 4)


@dataclass
class ShardResult:
    #: mean over all batches of the per-batch light curves, float64[nBins], or float64[nObjects, nBins] for
    #: the per object response (None without histogram). With several pipelines: of the first one.
    lightCurve: np.ndarray | None
    #: total number of batches (tasks) over all ranks
    nBatches: int
    #: event counters summed over all ranks and batches (name -> int), if a statistic callback is attached
    counters: dict = field(default_factory=dict)
    #: hits recorded by this rank, one dict of arrays per local batch (HitRecorder only)
    hits: list = field(default_factory=list)
    #: batches executed by this rank
    localBatches: list = field(default_factory=list)
    #: per pipeline name: mean light curve over the tasks of that pipeline (multi pipeline runs)
    lightCurves: dict = field(default_factory=dict)
    #: per pipeline name: number of tasks executed over all ranks
    taskCounts: dict = field(default_factory=dict)


class ShardedRun:
    """
    Runs batches of one tracer, or tasks of several named tracers (config C4: pipelines that differ in their
    light source and share target / response / generator, examples/03 of the reference), sharded over the ranks
    of a process group.

    Parameters
    ----------
    tracer: Tracer | dict[str, Tracer]
        tracer(s) whose response is a histogram (``HistogramHitResponse``, ``KernelHistogramHitResponse``,
        ``ObjectHistogramHitResponse``) and / or a ``HitRecorder``
    offset0: int
        Philox offset of batch 0
    group: torch.distributed.ProcessGroup | None
        process group; ``None`` = default group, or a single process if torch.distributed has not been initialised
    launcher: Callable[[int], None] | dict[str, Callable] | None
        executes configuration ``i`` of the pipeline and waits for it. Defaults to the tracer's own pipeline, run
        asynchronously. (The CPU tests of this module substitute the test oracle here; the product always runs the
        CUDA pipeline.)
    rngStride: int
        see the module documentation; 4 (default) = disjoint counter ranges, 1 = the reference's batch offsets
    """

    def __init__(self, tracer, *, offset0: int = 0, group=None, launcher=None, rngStride: int = RNG_STRIDE) -> None:
        self.tracers = dict(tracer) if isinstance(tracer, dict) else {"": tracer}
        if not self.tracers:
            raise ValueError("no tracer given")
        self.tracer = next(iter(self.tracers.values()))
        self.offset0 = int(offset0)
        self.group = group
        self.rngStride = int(rngStride)
        if launcher is not None and not isinstance(launcher, dict):
            launcher = {name: launcher for name in self.tracers}
        self._launchers = launcher
        self._pipelines = {name: pl.Pipeline(t.collectStages()) for name, t in self.tracers.items()}
        for t in self.tracers.values():
            t.rng.autoAdvance = 0  # offsets are set per batch
        # one counter range per task whatever pipeline it belongs to
        self._nRNG = max(t.nRNGSamples for t in self.tracers.values())

    # ------------------------------------------------------------------ process group
    @property
    def distributed(self) -> bool:
        return dist.is_available() and dist.is_initialized()

    @property
    def rank(self) -> int:
        return dist.get_rank(self.group) if self.distributed else 0

    @property
    def world(self) -> int:
        return dist.get_world_size(self.group) if self.distributed else 1

    def _reduceDevice(self) -> torch.device:
        if self.distributed and dist.get_backend(self.group) == "nccl":
            return torch.device("cuda", torch.cuda.current_device())
        if self.tracer._device == "cuda":
            return torch.device("cuda", torch.cuda.current_device())
        return torch.device("cpu")

    def _allReduce(self, t: torch.Tensor) -> torch.Tensor:
        if self.distributed and self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    # ------------------------------------------------------------------ one batch
    def _curveShape(self, resp) -> tuple:
        nBins = getattr(resp, "nBins", 0)
        return (resp.nObjects, nBins) if hasattr(resp, "nObjects") else (max(nBins, 1),)

    def _launch(self, name: str, i: int):
        """runs configuration i; returns the event that marks its completion (None: it has completed)"""
        if self._launchers is not None:
            self._launchers[name](i)
            return None
        return self._pipelines[name].runAsync(i)

    def _accumulate(self, acc: torch.Tensor, resp, i: int) -> None:
        """adds the light curve of the batch just launched (configuration i) to acc, stream ordered on a GPU"""
        if self._launchers is None and acc.is_cuda:
            red = getattr(resp, "_reducer", None)
            if red is not None:  # reduced, normalised histogram of this batch: still on the device
                acc.add_(red.histOut.view(acc.shape))
            else:  # per object histogram: its device array was cleared; take the pinned copy back (no host sync)
                acc.add_(resp._buffers[i].to(acc.device, non_blocking=True).view(acc.shape))
        else:
            acc.add_(torch.from_numpy(np.asarray(resp.result(i), dtype=np.float64)).to(acc.device).view(acc.shape))

    @staticmethod
    def _handOver(pending, processFn) -> None:
        b, resp, i, done = pending
        if done is not None:
            done.synchronize()
        processFn(b, resp.result(i))

    def _sync(self) -> None:
        if self.tracer._device == "cuda":
            torch.cuda.current_stream().synchronize()

    # ------------------------------------------------------------------ run
    def run(self, nBatches: int, *, keepHits: bool = False, processFn=None) -> ShardResult:
        """``nBatches`` batches of the (first) tracer with its current parameters."""
        name = next(iter(self.tracers))
        return self.runTasks([(name, {})] * nBatches, keepHits=keepHits, processFn=processFn)

    def runTasks(self, tasks: Iterable, *, keepHits: bool = False, processFn=None) -> ShardResult:
        """
        Tasks ``(pipeline name, {"stage__param": value, ...})`` - or plain parameter dicts with a single tracer -
        dealt to the ranks round robin (whole tasks: SURVEY.md 8e). Task ``t`` runs at the Philox offset of batch ``t``.

        ``processFn(task index, result of the task's response)``, if given, is called on the rank that ran the task once
        its result has reached the host (the per batch hook of the reference's ``PipelineScheduler``, pipeline.py:
        ``processFn(config, batch, args)``); the light curves are still accumulated on the device and reduced once.
        """
        tasks = [t if isinstance(t, tuple) else (next(iter(self.tracers)), t) for t in tasks]
        for name, _ in tasks:
            if name not in self.tracers:
                raise KeyError(f"unknown pipeline '{name}'")
        nTasks = len(tasks)
        dev = self._reduceDevice()
        names = list(self.tracers)
        shapes = [self._curveShape(self.tracers[n].response) for n in names]
        sizes = [int(np.prod(s)) for s in shapes]
        flat = torch.zeros(sum(sizes), dtype=torch.float64, device=dev)  # all light curves: ONE all-reduce
        accs, at = {}, 0
        for n, s, k in zip(names, shapes, sizes):
            accs[n] = flat[at:at + k].view(s)
            at += k
        res = ShardResult(None, nTasks)
        callbacks = []
        for t in self.tracers.values():
            cb = t.callback
            if cb is not None and all(cb is not c for c in callbacks):
                callbacks.append(cb)
                if hasattr(cb, "reset"):
                    cb.reset()
        mine, n = set(batchesOf(self.rank, self.world, nTasks)), 0
        pending = None
        for b, (name, params) in enumerate(tasks):
            # parameters persist from task to task like in a PipelineScheduler: every rank applies all of them (host
            # side only) and runs its own tasks, so the result does not depend on the number of ranks
            if params:
                self._pipelines[name].setParams(**params)
            if b not in mine:
                continue
            tr = self.tracers[name]
            i = n & 1
            n += 1
            tr.rng.setParam("offset", batchOffset(self.offset0, b, self._nRNG, self.rngStride))
            done = self._launch(name, i)
            res.localBatches.append(b)
            resp = tr.response
            if getattr(resp, "nBins", 0):
                self._accumulate(accs[name], resp, i)
            if processFn is not None:
                # the batch before this one is handed over now, while this one runs (the two configurations of the
                # pipeline keep their results apart), so the device does not wait for the host between batches
                if pending is not None:
                    self._handOver(pending, processFn)
                pending = (b, resp, i, done)
                if done is None:
                    self._handOver(pending, processFn)
                    pending = None
            if keepHits and hasattr(resp, "queue"):
                self._sync()
                v = resp.queue.view(i)
                count = min(int(v.count), resp.capacity)
                res.hits.append({f: np.array(v[f][:count]) for f in v.fields})
        if pending is not None:
            self._handOver(pending, processFn)
        self._sync()
        # the one collective of the run: light curves (+ event counters)
        self._allReduce(flat)
        counts = {n: sum(1 for t in tasks if t[0] == n) for n in names}
        res.taskCounts = counts
        for n in names:
            if getattr(self.tracers[n].response, "nBins", 0):
                res.lightCurves[n] = (accs[n] / max(counts[n], 1)).cpu().numpy()
        res.lightCurve = res.lightCurves.get(names[0])
        stat = [cb for cb in callbacks if getattr(cb, "Statistic", None) is not None]
        if stat:
            fields = [f[0] for f in stat[0].Statistic._fields_]
            c = torch.zeros(len(fields), dtype=torch.int64, device=dev)
            for cb in stat:
                c += torch.tensor([int(getattr(cb, f)) for f in fields], dtype=torch.int64, device=dev)
            self._allReduce(c)
            res.counters = {f: int(v) for f, v in zip(fields, c.cpu().tolist())}
        return res

    # ------------------------------------------------------------------ hits
    def gatherHits(self, res: ShardResult) -> dict:
        """
        All-gather of the hits a ``run(..., keepHits=True)`` recorded: first the per batch counts, then the packed
        arrays; every rank returns the same dict of arrays, hits concatenated in batch order (plus ``"batch"``, the
        batch each hit belongs to). One exchange of counts and one of data.
        """
        local = res.hits
        fields = list(local[0].keys()) if local else []
        world, dev = self.world, self._reduceDevice()
        # ---- counts per batch (fixed size: nBatches)
        counts = torch.zeros(res.nBatches, dtype=torch.int64, device=dev)
        for b, h in zip(res.localBatches, local):
            counts[b] = len(next(iter(h.values()))) if h else 0
        self._allReduce(counts)
        counts = counts.cpu().numpy()
        # ---- field layout from the first rank that has any hit (all ranks record the same item type)
        widths = {f: (int(np.prod(local[0][f].shape[1:])) if local[0][f].ndim > 1 else 1) for f in fields}
        meta = [None] * world
        if self.distributed and world > 1:
            dist.all_gather_object(meta, [(f, widths[f], str(local[0][f].dtype)) for f in fields], group=self.group)
        else:
            meta = [[(f, widths[f], str(local[0][f].dtype)) for f in fields]]
        layout = next((m for m in meta if m), [])
        if not layout:
            return {"batch": np.zeros(0, np.int64)}
        nCols = sum(w for _, w, _ in layout)
        perRank = [int(sum(counts[b] for b in batchesOf(r, world, res.nBatches))) for r in range(world)]
        rows = max(max(perRank), 1)
        # ---- packed payload: 32 bit words, rows padded to the largest rank
        mine = np.zeros((rows, nCols), np.uint32)
        at = 0
        for h in local:
            n = len(next(iter(h.values()))) if h else 0
            col = 0
            for f, w, _ in layout:
                mine[at:at + n, col:col + w] = np.ascontiguousarray(h[f]).reshape(n, w).view(np.uint32)
                col += w
            at += n
        payload = torch.from_numpy(mine.view(np.int32)).to(dev)
        if self.distributed and world > 1:
            parts = [torch.empty_like(payload) for _ in range(world)]
            dist.all_gather(parts, payload, group=self.group)
        else:
            parts = [payload]
        parts = [p.cpu().numpy().view(np.uint32) for p in parts]
        # ---- batch order
        cursor = [0] * world
        blocks, batchIds = [], []
        for b in range(res.nBatches):
            r, n = b % world, int(counts[b])
            blocks.append(parts[r][cursor[r]:cursor[r] + n])
            batchIds.append(np.full(n, b, np.int64))
            cursor[r] += n
        data = np.concatenate(blocks, 0) if blocks else np.zeros((0, nCols), np.uint32)
        out, col = {"batch": np.concatenate(batchIds) if batchIds else np.zeros(0, np.int64)}, 0
        for f, w, dt in layout:
            a = np.ascontiguousarray(data[:, col:col + w]).view(np.dtype(dt))
            out[f] = a[:, 0] if w == 1 else a
            col += w
        return out
