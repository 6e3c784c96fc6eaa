"""
The C-ABI boundary (include/theia_b200.h) without a GPU: the library loads, exports every
symbol the header declares, reports errors instead of computing when no device is present,
and the ctypes mirrors (theia_b200/_abi.py) have the layout a C compiler gives the header.
"""

import ctypes
import os
import re
import subprocess

import pytest

from theia_b200 import _abi, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "theia_b200.h")

MIRRORS = {
    "theia_philox_params_t": _abi.PhiloxParams,
    "theia_histogram_params_t": _abi.HistogramParams,
    "theia_geometry_t": _abi.Geometry,
    "theia_component_t": _abi.Component,
    "theia_response_t": _abi.Response,
    "theia_callback_t": _abi.Callback,
    "theia_trace_launch_t": _abi.TraceLaunch,
    "theia_as_sizes_t": _abi.ASSizes,
    "theia_instance_desc_t": _abi.InstanceDesc,
}


def declaredFunctions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    return sorted(set(re.findall(r"\b(theia_[a-z0-9_]+)\s*\(", text)))


def test_header_functions_are_bound():
    names = declaredFunctions()
    assert len(names) >= 17
    assert sorted(_lib.SYMBOLS) == names


def test_library_exports_every_symbol():
    lib = _lib.load()
    for name in declaredFunctions():
        assert hasattr(lib, name), name
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.libraryPath()], capture_output=True, text=True, check=True)
    exported = {ln.split()[-1] for ln in out.stdout.splitlines() if " T " in ln}
    assert set(declaredFunctions()) <= exported
    assert lib.theia_abi_version() == _abi.ABI_VERSION


def test_library_is_built_for_sm100a_only():
    out = subprocess.run(["cuobjdump", "-lelf", _lib.libraryPath()], capture_output=True, text=True, check=True)
    archs = set(re.findall(r"sm_(\d+a?)", out.stdout))
    assert archs == {"100a"}, archs


def test_header_compiles_as_c_and_layout_matches(tmp_path):
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', "int main(void) {"]
    for cname, mirror in MIRRORS.items():
        lines.append(f'  printf("{cname} size %zu\\n", sizeof({cname}));')
        for field in mirror._fields_:
            fname = field[0]
            lines.append(f'  printf("{cname} {fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "probe.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "probe"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-o", str(exe), str(src)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout
    seen = 0
    for ln in out.splitlines():
        cname, what, value = ln.split()
        mirror = MIRRORS[cname]
        if what == "size":
            assert ctypes.sizeof(mirror) == int(value), cname
        else:
            assert getattr(mirror, what).offset == int(value), (cname, what)
        seen += 1
    assert seen > 60


def test_no_device_is_an_error_not_a_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    lib = _lib.load()
    sm, major, minor = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    rc = lib.theia_device_info(ctypes.byref(sm), ctypes.byref(major), ctypes.byref(minor))
    assert rc != 0
    assert lib.theia_last_error()
    L = _abi.TraceLaunch()
    assert lib.theia_trace_forward(ctypes.byref(L)) != 0
    with pytest.raises(RuntimeError):
        _lib.check(lib.theia_trace_forward(ctypes.byref(L)))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "theia_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", text, flags=re.M), f
                assert "liboracle" not in text, f


def test_missing_library_raises(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "_LIB_PATH", str(tmp_path / "nothing.so"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.load()
