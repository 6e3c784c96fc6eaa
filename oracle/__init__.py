"""
ORACLE - TEST INFRASTRUCTURE ONLY.

ctypes binding of the CPU oracle (oracle/oracle.cpp). Only ``tests/``,
``__graft_entry__.smoke()`` and the CPU baseline / reference arm of ``bench.py`` may import
this package; the product (``theia_b200``) never does.

The oracle consumes the same C-ABI launch description as the CUDA library
(include/theia_b200.h) with host pointers in place of device pointers. The parity harness
therefore builds every object twice: once on the "cuda" device for the product and once
under ``theia_b200._device.useDevice("cpu")`` for the oracle.
"""

from __future__ import annotations

import ctypes
import os
import subprocess
from ctypes import POINTER, byref, c_float, c_int, c_int32, c_uint32, c_uint64, c_void_p

import numpy as np

from theia_b200 import _abi

_DIR = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_DIR, "_build", "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compiles the oracle with the Makefile next to this file."""
    if force or not os.path.exists(_SO):
        subprocess.run(["make", "-C", _DIR] + (["-B"] if force else []), check=True, capture_output=True)
    return _SO


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        l = ctypes.CDLL(_SO)
        l.orc_scene_create.restype = c_void_p
        l.orc_scene_create.argtypes = [POINTER(_abi.InstanceDesc), c_uint32, c_void_p, c_void_p, c_int]
        l.orc_scene_destroy.argtypes = [c_void_p]
        l.orc_trace_forward.argtypes = [POINTER(_abi.TraceLaunch), c_int, c_void_p]
        l.orc_philox_fill.argtypes = [POINTER(_abi.PhiloxParams), c_uint32, c_uint32, c_uint32, c_uint32, c_uint32,
                                      c_void_p]
        l.orc_philox_raw.argtypes = [c_void_p, c_void_p, c_uint32, c_void_p]
        l.orc_math_eval.argtypes = [c_uint32, c_void_p, c_void_p, c_uint32, c_void_p]
        l.orc_tlas_query.argtypes = [c_void_p, c_void_p, c_uint32, c_void_p, c_void_p, c_void_p]
        l.orc_hist_reduce.argtypes = [c_void_p, c_uint32, c_uint32, c_float, c_void_p, c_int]
        for fn in ("orc_target_sample", "orc_guide_sample"):
            getattr(l, fn).argtypes = [POINTER(_abi.Component), POINTER(_abi.PhiloxParams), c_void_p, c_uint32, c_void_p]
        for fn in ("orc_target_intersect", "orc_guide_eval"):
            getattr(l, fn).argtypes = [POINTER(_abi.Component), c_void_p, c_void_p, c_uint32, c_void_p]
        l.orc_hit_replay.argtypes = [c_void_p, c_uint32, c_uint32, POINTER(_abi.Response), c_uint32]
        l.orc_hit_replay_rng.argtypes = [c_void_p, c_uint32, c_uint32, POINTER(_abi.Response), c_void_p, c_uint32]
        _lib = l
    return _lib


def _ptr(a: np.ndarray) -> int:
    return a.ctypes.data


def philoxParams(key: int, offset: int) -> _abi.PhiloxParams:
    p = _abi.PhiloxParams()
    p.key[0], p.key[1] = key & 0xFFFFFFFF, (key >> 32) & 0xFFFFFFFF
    for k in range(4):
        p.offset[k] = (offset >> (32 * k)) & 0xFFFFFFFF
    return p


def philoxFill(key: int, offset: int, streams: int, samples: int, *, baseStream: int = 0, baseCount: int = 0,
               dim: int = 1) -> np.ndarray:
    out = np.zeros(streams * samples * dim, dtype=np.float32)
    p = philoxParams(key, offset)
    lib().orc_philox_fill(byref(p), baseStream, baseCount, streams, samples, dim, _ptr(out))
    return out.reshape(streams, samples, dim) if dim > 1 else out.reshape(streams, samples)


def philoxRaw(ctr: np.ndarray, key: np.ndarray) -> np.ndarray:
    ctr = np.ascontiguousarray(ctr, dtype=np.uint32).reshape(-1, 4)
    key = np.ascontiguousarray(key, dtype=np.uint32).reshape(-1, 2)
    out = np.zeros_like(ctr)
    lib().orc_philox_raw(_ptr(ctr), _ptr(key), len(ctr), _ptr(out))
    return out


MATH_FN = {"sin": 0, "cos": 1, "exp": 2, "log": 3, "pow": 4, "tanh": 5, "atan": 6}


def mathEval(fn: str, x: np.ndarray, y: np.ndarray | None = None) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.zeros_like(x) if y is None else np.ascontiguousarray(y, dtype=np.float32)
    out = np.zeros_like(x)
    lib().orc_math_eval(MATH_FN[fn], _ptr(x), _ptr(y), x.size, _ptr(out))
    return out


class OracleScene:
    """Oracle side acceleration structure of a ``theia_b200.scene.Scene`` built on the cpu device."""

    def __init__(self, scene, *, bruteForce: bool = False) -> None:
        if scene._device != "cpu":
            raise ValueError("the oracle needs a scene built under useDevice('cpu')")
        self._scene = scene  # keeps host geometry alive
        n = len(scene.instances)
        self.handle = lib().orc_scene_create(scene._descs, n, _ptr(scene._vertexCounts), _ptr(scene._triangleCounts),
                                             int(bruteForce))

    def query(self, rays: np.ndarray, *, stats: bool = False):
        """rays: float32[n, 8] = origin, tmax, direction, unused -> (ids int32[n,4], tuv float32[n,4])"""
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = len(rays)
        ids = np.zeros((n, 4), dtype=np.int32)
        tuv = np.zeros((n, 4), dtype=np.float32)
        st = np.zeros(4, dtype=np.uint64)
        lib().orc_tlas_query(self.handle, _ptr(rays), n, _ptr(ids), _ptr(tuv), _ptr(st) if stats else None)
        return (ids, tuv, st) if stats else (ids, tuv)

    def __del__(self) -> None:
        try:
            lib().orc_scene_destroy(self.handle)
        except Exception:
            pass


def trace(tracer, i: int = 0, *, scene: OracleScene | None = None, threads: int = 0, update: bool = True):
    """
    Runs one batch of ``tracer`` (built on the cpu device) on the oracle, writing into the
    tracer's own (host) response / callback buffers.
    Returns counters: uint64[6] = segments, responses, nodes, triangles, instances, rays.
    """
    if tracer._device != "cpu":
        raise ValueError("the oracle needs a tracer built under useDevice('cpu')")
    if update:
        for _, stage in tracer.collectStages():
            stage.update(i)
    L, keep = tracer._launch(i)
    if L.tracer in (_abi.TRACER_SCENE_FORWARD, _abi.TRACER_SCENE_PHOTON):
        if scene is None:
            scene = OracleScene(tracer.scene)
        L.tlas = scene.handle
    counters = np.zeros(6, dtype=np.uint64)
    rc = lib().orc_trace_forward(byref(L), threads, _ptr(counters))
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}")
    return counters


def histReduce(partials: np.ndarray, norm: float, *, clear: bool = False) -> np.ndarray:
    partials = np.ascontiguousarray(partials, dtype=np.float32)
    nHist, nBins = partials.shape
    out = np.zeros(nBins, dtype=np.float32)
    lib().orc_hist_reduce(_ptr(partials), nHist, nBins, norm, _ptr(out), int(clear))
    return out


def _componentOf(stage):
    """(Component, keep-alive) of a target / target guide stage with its current parameters"""
    from theia_b200.trace import _component
    stage.update(0)
    keep: list = []
    return _component(stage, 0, keep), keep


_TARGET_FIELDS = [("position", 0, 3), ("normal", 3, 6), ("dist", 6, 7), ("objPosition", 7, 10), ("objNormal", 10, 13),
                  ("prob", 13, 14), ("valid", 14, 15)]


def _unpack(out: np.ndarray, fields) -> dict:
    return {name: (out[:, a] if b - a == 1 else out[:, a:b]) for name, a, b in fields}


def targetSample(target, observer: np.ndarray, *, key: int = 1, offset: int = 0) -> dict:
    """sampleTarget(observer[i], idx=i, dim=0) of a Target stage"""
    comp, keep = _componentOf(target)
    obs = np.ascontiguousarray(observer, dtype=np.float32)
    out = np.zeros((len(obs), 15), np.float32)
    lib().orc_target_sample(ctypes.byref(comp), ctypes.byref(philoxParams(key, offset)), _ptr(obs), len(obs), _ptr(out))
    return _unpack(out, _TARGET_FIELDS)


def targetIntersect(target, observer: np.ndarray, direction: np.ndarray) -> dict:
    comp, keep = _componentOf(target)
    obs, d = np.ascontiguousarray(observer, dtype=np.float32), np.ascontiguousarray(direction, dtype=np.float32)
    out = np.zeros((len(obs), 15), np.float32)
    lib().orc_target_intersect(ctypes.byref(comp), _ptr(obs), _ptr(d), len(obs), _ptr(out))
    return _unpack(out, _TARGET_FIELDS)


_GUIDE_FIELDS = [("dir", 0, 3), ("dist", 3, 4), ("prob", 4, 5)]


def guideSample(guide, observer: np.ndarray, *, key: int = 1, offset: int = 0) -> dict:
    comp, keep = _componentOf(guide)
    obs = np.ascontiguousarray(observer, dtype=np.float32)
    out = np.zeros((len(obs), 5), np.float32)
    lib().orc_guide_sample(ctypes.byref(comp), ctypes.byref(philoxParams(key, offset)), _ptr(obs), len(obs), _ptr(out))
    return _unpack(out, _GUIDE_FIELDS)


def guideEval(guide, observer: np.ndarray, direction: np.ndarray) -> dict:
    comp, keep = _componentOf(guide)
    obs, d = np.ascontiguousarray(observer, dtype=np.float32), np.ascontiguousarray(direction, dtype=np.float32)
    out = np.zeros((len(obs), 5), np.float32)
    lib().orc_guide_eval(ctypes.byref(comp), _ptr(obs), _ptr(d), len(obs), _ptr(out))
    return _unpack(out, _GUIDE_FIELDS)


# ------------------------------------------------------------------ unit seam (mirror of theia_unit_eval)
def evalUnit(unit: int, rows: np.ndarray | None, n: int | None = None, **kw) -> np.ndarray:
    """
    The oracle's evaluation of one unit (oracle/units.h): same arguments as
    ``theia_b200._units.evalUnit``, components built under ``useDevice("cpu")``.
    """
    from theia_b200 import _units
    l = lib()
    l.orc_unit_eval.argtypes = [POINTER(_abi.UnitArgs)]
    wi, wo = _units.WIDTHS[unit]
    if rows is not None:
        rows = np.ascontiguousarray(rows, dtype=np.float32).reshape(-1, wi)
        n = len(rows)
    A = _units.unitArgs(unit, int(n), **kw)
    out = np.zeros((int(n), wo), dtype=np.float32)
    A.in_ = _ptr(rows) if rows is not None else None
    A.out = _ptr(out)
    rc = l.orc_unit_eval(byref(A))
    if rc != 0:
        raise RuntimeError(f"oracle unit error {rc}")
    return out


def runSampler(sampler, i: int = 0) -> None:
    """
    Runs a sampler stage (``light.LightSampler``, ``testing.TargetSampler``, ``testing.TargetGuideSampler``)
    built on the cpu device on the oracle: updates its stages, evaluates, and fills the host buffers the
    stage's accessors (``lightQueue.view(i)``, ``getResults(i)``) read.
    """
    l = lib()
    l.orc_unit_eval.argtypes = [POINTER(_abi.UnitArgs)]
    if sampler._device != "cpu":
        raise ValueError("the oracle needs a sampler built under useDevice('cpu')")
    stages = sampler.collectStages()
    for s in stages:
        s.update(i)
    for s in stages[:-1]:  # host sources copy their queues "to the device"
        for cmd in s.run(i):
            cmd()
    A = sampler._unitArgs(i)
    rc = l.orc_unit_eval(byref(A))
    if rc != 0:
        raise RuntimeError(f"oracle unit error {rc}")
    if A.unit == _abi.UNIT_LIGHT_SAMPLE:
        for q in (sampler.lightQueue, sampler.wavelengthQueue):
            for cmd in q.run(i):
                cmd()
    else:
        nbytes = 4 * sampler.batchSize * sampler._width
        sampler._buffer[i][:nbytes].view(sampler._tensor.dtype).copy_(sampler._tensor)


def runReplay(replay, i: int = 0) -> None:
    """
    Runs a ``response.HitReplay`` stage built on the cpu device on the oracle (sh/response.replay.glsl:35-50):
    feeds the hits of its host queue to its response and fills the response's host side results.
    """
    if replay._device != "cpu":
        raise ValueError("the oracle needs a replay stage built under useDevice('cpu')")
    stages = replay.collectStages()
    for s in stages:
        s.update(i)
    for cmd in replay.queue.run(i):  # host copy -> "device" copy of the queue
        cmd()
    nBlocks = -(replay.capacity // -replay.blockSize)
    r = replay.response._abiResponse(i, nBlocks)
    rng = replay.rng._abiParams(i) if replay.rng is not None else None
    rc = lib().orc_hit_replay_rng(replay.queue.tensor.data_ptr(), replay.capacity, int(replay.polarized), byref(r),
                                  ctypes.addressof(rng) if rng is not None else None, replay.blockSize)
    if rc != 0:
        raise RuntimeError(f"oracle replay error {rc}")
    resp = replay.response
    red = getattr(resp, "_reducer", None)
    if red is not None:
        p = red._block("Params", i)
        parts = red.histIn.numpy().reshape(-1, resp.nBins)
        red._buffer[i].numpy()[:] = histReduce(parts, p.normalization, clear=True)
        red.histIn.zero_()
    else:
        for cmd in resp.run(i):
            cmd()
