"""
Multi-GPU sharding of a tracer run: one process per GPU, paths are
independent, so the only exchange is one all-reduce of the accumulated light curve (and of
the event counters) at the end of the run.

Batch ``b`` of a run always uses the Philox offset ``offset0 + b * nRNGSamples`` and the
streams ``0 .. batchSize-1`` - the same counters a single-GPU run (and a reference run with
``autoAdvance = nRNGSamples``, src/theia/random.py:278-282) uses for its b-th batch - and
rank ``r`` of ``R`` executes the batches ``{b : b mod R = r}``. Results therefore do not
depend on the number of GPUs except for the float summation order of the final reduction,
which is done in float64.

The collective runs on ``torch.distributed`` (NCCL over NVLink on the GPU box; gloo in the
CPU tests of the host logic). Nothing here computes: batches are executed by the pipeline
handed in, i.e. by the CUDA library.
"""

from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable

import numpy as np
import torch
import torch.distributed as dist

from theia_b200 import pipeline as pl

__all__ = ["batchesOf", "batchOffset", "ShardResult", "ShardedRun"]


def batchesOf(rank: int, world: int, nBatches: int) -> range:
    """Batches rank ``rank`` of ``world`` executes: dealt round robin."""
    if not 0 <= rank < world:
        raise ValueError("rank must be in [0, world)")
    return range(rank, nBatches, world)


def batchOffset(offset0: int, batch: int, nRNGSamples: int) -> int:
    """Philox offset of a batch: 128 bit counter arithmetic wraps like the device's."""
    return (offset0 + batch * nRNGSamples) & ((1 << 128) - 1)


@dataclass
class ShardResult:
    #: mean over all batches of the per-batch light curves, float64[nBins], or float64[nObjects, nBins] for
    #: the per object response (None without histogram)
    lightCurve: np.ndarray | None
    #: total number of batches over all ranks
    nBatches: int
    #: event counters summed over all ranks and batches (name -> int), if a statistic callback is attached
    counters: dict = field(default_factory=dict)
    #: hits recorded by this rank, one dict of arrays per local batch (HitRecorder only)
    hits: list = field(default_factory=list)
    #: batches executed by this rank
    localBatches: list = field(default_factory=list)


class ShardedRun:
    """
    Runs ``nBatches`` batches of a tracer sharded over the ranks of a process group.

    Parameters
    ----------
    tracer: theia_b200.trace.Tracer
        tracer whose response is a histogram (``HistogramHitResponse``, ``KernelHistogramHitResponse``,
        ``ObjectHistogramHitResponse``) and / or a ``HitRecorder``
    offset0: int
        Philox offset of batch 0
    group: torch.distributed.ProcessGroup | None
        process group; ``None`` = default group, or a single process if torch.distributed
        has not been initialised
    launcher: Callable[[int], None] | None
        executes configuration ``i`` of the pipeline and waits for it. Defaults to
        ``Pipeline(tracer.collectStages()).run``. (The CPU tests of this module substitute the
        test oracle here; the product always runs the CUDA pipeline.)
    """

    def __init__(self, tracer, *, offset0: int = 0, group=None, launcher: Callable[[int], None] | None = None) -> None:
        self.tracer = tracer
        self.offset0 = int(offset0)
        self.group = group
        if launcher is None:
            self._pipeline = pl.Pipeline(tracer.collectStages())
            launcher = self._pipeline.run
        self._launcher = launcher
        self.tracer.rng.autoAdvance = 0  # offsets are set per batch

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
        return torch.device("cpu")

    def _allReduce(self, t: torch.Tensor) -> torch.Tensor:
        if self.distributed and self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    # ------------------------------------------------------------------ run
    def run(self, nBatches: int, *, keepHits: bool = False) -> ShardResult:
        tr = self.tracer
        resp, cb = tr.response, tr.callback
        nRNG = tr.nRNGSamples
        dev = self._reduceDevice()
        nBins = getattr(resp, "nBins", 0)
        shape = (resp.nObjects, nBins) if hasattr(resp, "nObjects") else (max(nBins, 1),)
        acc = torch.zeros(shape, dtype=torch.float64, device=dev)
        res = ShardResult(None, nBatches)
        stat = getattr(cb, "Statistic", None)  # EventStatisticCallback
        counterNames = [f[0] for f in stat._fields_] if stat is not None else []
        if cb is not None and hasattr(cb, "reset"):
            cb.reset()
        for n, b in enumerate(batchesOf(self.rank, self.world, nBatches)):
            i = n & 1
            tr.rng.setParam("offset", batchOffset(self.offset0, b, nRNG))
            self._launcher(i)
            res.localBatches.append(b)
            if nBins:
                h = resp.result(i)
                acc += torch.from_numpy(np.asarray(h, dtype=np.float64)).to(dev)
            if keepHits and hasattr(resp, "queue"):
                v = resp.queue.view(i)
                count = min(int(v.count), resp.capacity)
                res.hits.append({name: np.array(v[name][:count]) for name in v.fields})
        # the one collective of the run: light curve (+ event counters)
        self._allReduce(acc)
        if nBins:
            res.lightCurve = (acc / max(nBatches, 1)).cpu().numpy()
        if counterNames:
            c = torch.tensor([int(getattr(cb, name)) for name in counterNames], dtype=torch.int64, device=dev)
            self._allReduce(c)
            res.counters = {name: int(v) for name, v in zip(counterNames, c.cpu().tolist())}
        return res
