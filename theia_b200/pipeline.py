"""
Pipeline stages, pipelines and schedulers: the in-repo stand-in for the part of
``hephaistos.pipeline`` that theia's user code touches (docs/pipeline/pipeline.md:1-94,
docs/pipeline/extend.md:216-346 of the reference).

Differences to the Vulkan original, by design:
 * a "command" is a Python callable that enqueues work on the current CUDA stream
   (kernel launches through the C-ABI, async copies); ``Pipeline`` collects them once
   per configuration, exactly like the reference collects command buffers.
 * parameter blocks are plain ``ctypes.Structure`` instances kept three times: the live
   host state and one snapshot per pipeline configuration (0/1), taken in ``update(i)``.
"""

from __future__ import annotations

import ctypes
import queue as _queue
import threading
from ctypes import Structure
from typing import Any, Callable, Iterable

import numpy as np
import torch

from theia_b200 import _device
from theia_b200.glsl import _Vec

__all__ = [
    "Command",
    "Pipeline",
    "PipelineParams",
    "PipelineScheduler",
    "PipelineStage",
    "RetrieveTensorStage",
    "SourceCodeMixin",
    "runPipeline",
]

Command = Callable[[], None]
PipelineParams = dict


def _toPython(value):
    """ctypes field value -> plain python value"""
    if isinstance(value, _Vec):
        return tuple(value)
    if isinstance(value, ctypes.Array):
        return tuple(_toPython(v) for v in value)
    if isinstance(value, Structure):
        if hasattr(type(value), "value"):
            return value.value
        return value
    return value


def _assign(struct: Structure, field: str, ctype, value) -> None:
    """assigns a plain python value to a ctypes field"""
    if issubclass(ctype, _Vec):
        getattr(struct, field).assign(value)
    elif issubclass(ctype, ctypes.Array):
        arr = getattr(struct, field)
        flat = np.asarray(value).reshape(-1)
        if len(flat) != len(arr):
            raise ValueError(f"expected {len(arr)} values for {field}")
        for k, v in enumerate(flat):
            arr[k] = v
    elif issubclass(ctype, Structure):
        if hasattr(ctype, "value") and not isinstance(value, ctype):
            getattr(struct, field).value = value
        elif isinstance(value, ctype):
            ctypes.memmove(
                ctypes.addressof(getattr(struct, field)), ctypes.addressof(value), ctypes.sizeof(ctype)
            )
        else:
            flat = np.asarray(value, dtype=np.float32).reshape(-1)
            dst = getattr(struct, field)
            if flat.nbytes != ctypes.sizeof(ctype):
                raise ValueError(f"cannot assign value to {field}")
            ctypes.memmove(ctypes.addressof(dst), flat.ctypes.data, flat.nbytes)
    else:
        setattr(struct, field, value)


class PipelineStage:
    """
    Smallest unit of a pipeline (docs/pipeline/pipeline.md:24-55). Owns named parameter
    blocks mirrored as ctypes structures. Fields are readable/writable as attributes and via
    ``getParam/setParam/getParams/setParams``; names starting with ``_`` are private (not
    listed, not attributes, but settable through ``setParams``); ``extra`` names Python-only
    properties that are exposed through the same interface.
    """

    name = "stage"

    def __init__(self, params: dict[str, type[Structure]] = {}, extra: set[str] = set()) -> None:
        object.__setattr__(self, "_fields", {})
        object.__setattr__(self, "_extra", set(extra))
        object.__setattr__(self, "_local", {b: t() for b, t in params.items()})
        object.__setattr__(self, "_configs", [{b: t() for b, t in params.items()} for _ in range(2)])
        for block, t in params.items():
            for fname, ftype in t._fields_:
                self._fields[fname] = (block, ftype)
        object.__setattr__(self, "_device", _device.currentDevice())

    # ---- attribute sugar
    def __getattr__(self, name: str):
        fields = object.__getattribute__(self, "__dict__").get("_fields")
        if fields is not None and name in fields and not name.startswith("_"):
            return self.getParam(name)
        raise AttributeError(f"{type(self).__name__!r} object has no attribute {name!r}")

    def __setattr__(self, name: str, value) -> None:
        fields = self.__dict__.get("_fields")
        if (
            fields is not None
            and name in fields
            and not name.startswith("_")
            and not isinstance(getattr(type(self), name, None), property)
        ):
            self.setParam(name, value)
        else:
            object.__setattr__(self, name, value)

    # ---- parameter interface
    @property
    def fields(self) -> set[str]:
        """Names of all public parameters"""
        return {f for f in self._fields if not f.startswith("_")} | self._extra

    def getParam(self, name: str) -> Any:
        if name in self._fields:
            block, _ = self._fields[name]
            return _toPython(getattr(self._local[block], name))
        if name in self._extra:
            return getattr(self, name)
        raise KeyError(name)

    def setParam(self, name: str, value: Any) -> None:
        if name in self._fields:
            block, ftype = self._fields[name]
            _assign(self._local[block], name, ftype, value)
        elif name in self._extra:
            setattr(self, name, value)
        # unknown parameters are ignored like in hephaistos

    def getParams(self) -> dict[str, Any]:
        return {n: self.getParam(n) for n in self.fields}

    def setParams(self, **kwargs) -> None:
        for k, v in kwargs.items():
            self.setParam(k, v)

    # ---- pipeline hooks
    def _finishParams(self, i: int) -> None:
        """Hook called before the parameters are committed to configuration i"""
        pass

    def update(self, i: int) -> None:
        """Commits the current parameters to configuration i"""
        self._finishParams(i)
        for block, src in self._local.items():
            dst = self._configs[i][block]
            ctypes.memmove(ctypes.addressof(dst), ctypes.addressof(src), ctypes.sizeof(src))

    def _block(self, name: str, i: int) -> Structure:
        """Parameter block as committed to configuration i"""
        return self._configs[i][name]

    def _firstBlock(self, i: int) -> Structure | None:
        for b in self._configs[i].values():
            return b
        return None

    def run(self, i: int) -> list[Command]:
        """Commands this stage contributes to configuration i"""
        return []


class SourceCodeMixin(PipelineStage):
    """
    Base of all components that are plugged into a tracer (docs/pipeline/extend.md:216-225).
    The reference returns GLSL here; this build selects pre-compiled CUDA variants, so
    ``sourceCode`` is a short variant description kept for API compatibility.
    """

    name = "component"
    #: C-ABI kind of the component (enum values of include/theia_b200.h)
    _kind: int = 0

    def __init__(self, params: dict[str, type[Structure]] = {}, extra: set[str] = set()) -> None:
        super().__init__(params, extra)

    @property
    def _flags(self) -> int:
        return 0

    @property
    def sourceCode(self) -> str:
        return f"// cuda variant: {type(self).__name__} kind={self._kind} flags={self._flags}\n"

    def bindParams(self, program, i: int) -> None:  # kept for API compatibility
        pass


class RetrieveTensorStage(PipelineStage):
    """Copies a device tensor into one of two host buffers (hephaistos.pipeline.RetrieveTensorStage)."""

    name = "retrieve"

    def __init__(self, tensor: torch.Tensor) -> None:
        super().__init__()
        self._tensor = tensor
        self._buffers = [
            torch.zeros(tensor.shape, dtype=tensor.dtype, pin_memory=torch.cuda.is_available()) for _ in range(2)
        ]

    def buffer(self, i: int) -> torch.Tensor:
        return self._buffers[i]

    def view(self, i: int) -> np.ndarray:
        return self._buffers[i].numpy()

    def run(self, i: int) -> list[Command]:
        return [lambda: self._buffers[i].copy_(self._tensor, non_blocking=True)]


def _nameStages(stages) -> list[tuple[str, PipelineStage]]:
    named = []
    for s in stages:
        if isinstance(s, tuple):
            named.append((s[0], s[1]))
        else:
            named.append((s.name, s))
    return named


class Pipeline:
    """
    Sequence of stages run in order (docs/pipeline/pipeline.md:44-64). Parameters of the
    stages are addressed as ``stage__parameter``.
    """

    def __init__(self, stages: Iterable[PipelineStage | tuple[str, PipelineStage]]) -> None:
        self._stages = _nameStages(stages)
        self._commands = [
            [cmd for _, s in self._stages for cmd in s.run(i)] for i in range(2)
        ]
        devs = {s._device for _, s in self._stages}
        self._device = "cuda" if "cuda" in devs else "cpu"

    @property
    def stages(self) -> list[tuple[str, PipelineStage]]:
        return self._stages

    def getParams(self) -> dict[str, Any]:
        return {f"{n}__{k}": v for n, s in self._stages for k, v in s.getParams().items()}

    def setParams(self, **kwargs) -> None:
        for key, value in kwargs.items():
            stage, sep, param = key.partition("__")
            if not sep:
                continue
            for n, s in self._stages:
                if n == stage:
                    s.setParam(param, value)

    def update(self, i: int) -> None:
        for _, s in self._stages:
            s.update(i)

    def runAsync(self, i: int, *, update: bool = True) -> "torch.cuda.Event | None":
        """Enqueues one run of configuration i; returns an event marking its completion."""
        if update:
            self.update(i)
        for cmd in self._commands[i]:
            cmd()
        if self._device == "cuda":
            ev = torch.cuda.Event()
            ev.record()
            return ev
        return None

    def run(self, i: int, *, update: bool = True) -> None:
        ev = self.runAsync(i, update=update)
        if ev is not None:
            ev.synchronize()


def runPipeline(stages: Iterable[PipelineStage | tuple[str, PipelineStage]], i: int = 0) -> None:
    """Creates a pipeline from the stages, runs it once and waits for it to finish."""
    Pipeline(stages).run(i)


class PipelineScheduler:
    """
    Orchestrates the submission of tasks onto a pipeline (docs/pipeline/pipeline.md:66-94):
    an update thread prepares configuration ``n % 2`` and enqueues the GPU work, a process
    thread waits for the GPU and calls ``processFn(config, batch, args)``. A configuration
    is only reused once its previous result has been processed.

    ``pipeline`` may also be a dict of named pipelines (examples/03_multiple_lightsources.ipynb
    cell 6); tasks are then ``(name, params)`` tuples. The pipelines may share stages (target,
    response, rng, ...): configurations alternate over all tasks, whichever pipeline runs them.
    """

    def __init__(
        self,
        pipeline: "Pipeline | dict[str, Pipeline]",
        *,
        queueSize: int = 0,
        processFn: Callable[[int, int, Any], None] | None = None,
        processArgs: Any = None,
    ) -> None:
        if isinstance(pipeline, dict):
            if not pipeline:
                raise ValueError("no pipeline given")
            self._pipelines = dict(pipeline)
            self._pipeline = next(iter(self._pipelines.values()))
        else:
            self._pipelines = None
            self._pipeline = pipeline
        self._processFn = processFn
        self._processArgs = processArgs
        self._tasks: _queue.Queue = _queue.Queue(queueSize)
        self._done: _queue.Queue = _queue.Queue()
        self._free = [threading.Semaphore(1), threading.Semaphore(1)]
        self._total = 0
        self._finished = 0
        self._submitted = 0
        self._cond = threading.Condition()
        self._error: BaseException | None = None
        self._alive = True
        devices = {p._device for p in (self._pipelines.values() if self._pipelines else [self._pipeline])}
        self._stream = torch.cuda.Stream() if "cuda" in devices else None
        self._tUpdate = threading.Thread(target=self._updateLoop, daemon=True)
        self._tProcess = threading.Thread(target=self._processLoop, daemon=True)
        self._tUpdate.start()
        self._tProcess.start()

    @property
    def pipeline(self) -> Pipeline:
        return self._pipeline

    @property
    def tasksFinished(self) -> int:
        return self._finished

    @property
    def totalTasks(self) -> int:
        return self._total

    def schedule(self, tasks: Iterable, *, timeout: float | None = None) -> None:
        """Queues tasks: parameter dicts, or ``(pipelineName, params)`` with named pipelines."""
        for t in tasks:
            if self._pipelines is not None:
                if not (isinstance(t, tuple) and len(t) == 2):
                    raise ValueError("with named pipelines a task is a tuple (name, params)")
                name, params = t
                if name not in self._pipelines:
                    raise KeyError(f"unknown pipeline '{name}'")
                item = (name, dict(params))
            else:
                item = (None, dict(t))
            with self._cond:
                self._total += 1
            self._tasks.put(item, timeout=timeout)

    def _fail(self, err: BaseException) -> None:
        with self._cond:
            if self._error is None:
                self._error = err
            self._cond.notify_all()

    def _updateLoop(self) -> None:
        while True:
            task = self._tasks.get()
            if task is None:
                self._done.put(None)
                return
            n = self._submitted
            i = n % 2
            self._free[i].acquire()
            ev = None
            try:
                if self._error is None:
                    name, params = task
                    pipeline = self._pipeline if name is None else self._pipelines[name]
                    pipeline.setParams(**params)
                    if self._stream is not None:
                        with torch.cuda.stream(self._stream):
                            ev = pipeline.runAsync(i)
                    else:
                        ev = pipeline.runAsync(i)
            except BaseException as err:  # surfaced by wait()
                self._fail(err)
            self._submitted += 1
            self._done.put((i, n, ev))

    def _processLoop(self) -> None:
        while True:
            item = self._done.get()
            if item is None:
                return
            i, n, ev = item
            try:
                if ev is not None:
                    ev.synchronize()
                if self._processFn is not None and self._error is None:
                    self._processFn(i, n, self._processArgs)
            except BaseException as err:
                self._fail(err)
            self._free[i].release()
            with self._cond:
                self._finished += 1
                self._cond.notify_all()

    def wait(self) -> None:
        """Blocks until all scheduled tasks have been processed; re-raises worker errors."""
        with self._cond:
            self._cond.wait_for(lambda: self._finished >= self._total)
            if self._error is not None:
                err, self._error = self._error, None
                raise err

    def destroy(self) -> None:
        if self._alive:
            self._alive = False
            self._tasks.put(None)
            self._tUpdate.join()
            self._tProcess.join()

    def __del__(self) -> None:
        try:
            self.destroy()
        except Exception:
            pass


class DynamicTask:
    """
    A task that decides, batch by batch, whether it needs more batches (the role of
    ``hephaistos.pipeline.DynamicTask`` in src/theia/task.py:22-196): the scheduler issues
    ``initialBatchCount`` batches with ``parameters`` on ``pipeline``; every finished batch is
    handed to ``processBatch(config)``, whose return value is the number of further batches to
    issue. Once nothing is in flight and nothing more was requested, ``onTaskFinished`` runs.
    """

    def __init__(self, parameters: dict, pipeline: str | None = None, *, initialBatchCount: int = 1) -> None:
        if initialBatchCount < 1:
            raise ValueError("initialBatchCount must be at least 1")
        self._parameters = dict(parameters)
        self._pipelineName = pipeline
        self._initialBatchCount = initialBatchCount
        self._inFlight = 0
        self._issued = 0
        self._finished = False

    @property
    def parameters(self) -> dict:
        return self._parameters

    @property
    def pipeline(self) -> str | None:
        return self._pipelineName

    @property
    def initialBatchCount(self) -> int:
        return self._initialBatchCount

    @property
    def batchesRemaining(self) -> int:
        """Batches issued for this task that have not been processed yet"""
        return self._inFlight

    @property
    def totalBatchesIssued(self) -> int:
        return self._issued

    @property
    def isFinished(self) -> bool:
        return self._finished

    def processBatch(self, config: int) -> int:
        """Consumes the result of configuration ``config``; returns how many more batches to issue."""
        return 0

    def onTaskFinished(self) -> None:
        pass


class DynamicTaskScheduler:
    """
    Runs ``DynamicTask``s on one pipeline or on named pipelines: batches of different tasks
    interleave on the two configurations exactly like plain tasks of a ``PipelineScheduler``.
    """

    def __init__(self, pipeline: "Pipeline | dict[str, Pipeline]", *, queueSize: int = 0) -> None:
        self._named = isinstance(pipeline, dict)
        self._owners: _queue.Queue = _queue.Queue()
        self._lock = threading.Lock()
        self._scheduler = PipelineScheduler(pipeline, queueSize=queueSize, processFn=self._process)

    def _issue(self, task: DynamicTask, n: int) -> None:
        if n <= 0:
            return
        with self._lock:
            task._inFlight += n
            task._issued += n
            for _ in range(n):
                self._owners.put(task)
        item = (task.pipeline, task.parameters) if self._named else task.parameters
        self._scheduler.schedule([item] * n)

    def _process(self, config: int, batch: int, args) -> None:
        task: DynamicTask = self._owners.get()  # batches finish in the order they were issued
        with self._lock:
            task._inFlight -= 1
        more = int(task.processBatch(config) or 0)
        if more > 0:
            self._issue(task, more)
        elif task._inFlight == 0 and not task._finished:
            task._finished = True
            task.onTaskFinished()

    def scheduleTask(self, task: DynamicTask) -> None:
        if self._named and task.pipeline is None:
            raise ValueError("the task must name its pipeline")
        self._issue(task, task.initialBatchCount)

    def scheduleTasks(self, tasks: Iterable[DynamicTask]) -> None:
        for t in tasks:
            self.scheduleTask(t)

    def wait(self) -> None:
        self._scheduler.wait()

    def destroy(self) -> None:
        self._scheduler.destroy()
