"""loads magicsoup_b200._lib (C++ / pybind11, csrc/lib_host.cpp), building it first if needed"""
from __future__ import annotations

import importlib

_mod = None


def lib():
    global _mod
    if _mod is None:
        from magicsoup_b200 import build as _build

        _build.build_host_lib()  # raises if g++ / pybind11 are missing: there is no Python fallback
        _mod = importlib.import_module("magicsoup_b200._lib")
    return _mod
