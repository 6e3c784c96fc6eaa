"""
Builds ``theia_b200/libtheia_b200.so`` (the C-ABI library of include/theia_b200.h) in-tree
with nvcc for sm_100a. Called by ``__graft_entry__.build()``; can also be run directly:
``python -m theia_b200.csrc.build [--force]``.

Flags that are part of the arithmetic contract: ``-fmad=false`` (no contraction;
fused multiply-adds are written explicitly), IEEE division and square root (nvcc defaults).
"""

from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

_DIR = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.dirname(_DIR)
_OUT = os.path.join(_PKG, "libtheia_b200.so")
_OBJ = os.path.join(_DIR, "_obj")

SOURCES = ["api.cu", "bvh_build.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC",
    "-Xcudafe", "--diag_suppress=177",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.sep not in cand or os.path.exists(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def _deps() -> list[str]:
    files = [os.path.join(_DIR, f) for f in os.listdir(_DIR) if f.endswith((".cu", ".cuh"))]
    files.append(os.path.join(os.path.dirname(_PKG), "include", "theia_b200.h"))
    return files


def upToDate() -> bool:
    if not os.path.exists(_OUT):
        return False
    t = os.path.getmtime(_OUT)
    return all(os.path.getmtime(f) <= t for f in _deps())


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and upToDate():
        return _OUT
    os.makedirs(_OBJ, exist_ok=True)
    nvcc = _nvcc()

    def compile_one(src: str) -> str:
        obj = os.path.join(_OBJ, src.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, "-c", os.path.join(_DIR, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(len(SOURCES)) as pool:
        objs = list(pool.map(compile_one, SOURCES))
    cmd = [nvcc, "-shared", "-o", _OUT, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart_static",
           "-Xcompiler", "-fPIC"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return _OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
