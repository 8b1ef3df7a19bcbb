"""
Builds magicsoup_b200/libmagicsoup_b200.so in-tree with nvcc for sm_100a only.
`python -m magicsoup_b200.build` or __graft_entry__.build().
"""
from __future__ import annotations

import shutil
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
LIB = HERE / "libmagicsoup_b200.so"
SOURCES = ("ms_api.cu", "kinetics.cu", "diffusion.cu", "params.cu", "peer.cu")
# -fmad=false: the spec order has no fused multiply-add (bit parity with the CPU oracle)
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-fmad=false",
    "-std=c++17", "-Xcompiler", "-fPIC", "--shared",
]


def nvcc() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(cand).exists():
        raise RuntimeError("nvcc not found")
    return cand


STAMP = HERE / "libmagicsoup_b200.so.srchash"


def source_hash() -> str:
    """content hash of everything the library is built from (mtimes do not survive a snapshot)"""
    import hashlib

    h = hashlib.sha256(" ".join(NVCC_FLAGS).encode())
    deps = sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cuh")) + [HERE.parent / "include" / "magicsoup_b200.h"]
    for d in deps:
        h.update(d.name.encode())
        h.update(d.read_bytes())
    return h.hexdigest()


def needs_build() -> bool:
    if not LIB.exists() or not STAMP.exists():
        return True
    return STAMP.read_text().strip() != source_hash()


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not needs_build():
        return LIB
    cmd = [nvcc(), *NVCC_FLAGS, *(["-Xptxas", "-v"] if verbose else []),
           "-o", str(LIB), *[str(CSRC / s) for s in SOURCES]]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    STAMP.write_text(source_hash())
    return LIB


# ---------------------------------------------------------------- host helpers (_lib, C++ / pybind11)
def lib_path() -> Path:
    import sysconfig

    return HERE / ("_lib" + (sysconfig.get_config_var("EXT_SUFFIX") or ".so"))


LIB_STAMP = HERE / "_lib.srchash"


def _lib_hash() -> str:
    import hashlib

    return hashlib.sha256((CSRC / "lib_host.cpp").read_bytes()).hexdigest()


def build_host_lib(force: bool = False) -> Path:
    """magicsoup_b200._lib: the interface of the reference's Rust extension, in C++ (g++, pybind11, OpenMP)"""
    import sys

    out = lib_path()
    if not force and out.exists() and LIB_STAMP.exists() and LIB_STAMP.read_text().strip() == _lib_hash():
        return out
    import pybind11
    import sysconfig

    cxx = shutil.which("g++")
    if cxx is None:
        raise RuntimeError("g++ not found")
    cmd = [cxx, "-O2", "-std=c++17", "-shared", "-fPIC", "-fopenmp", "-fvisibility=hidden",
           f"-I{pybind11.get_include()}", f"-I{sysconfig.get_paths()['include']}",
           str(CSRC / "lib_host.cpp"), "-o", str(out)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("g++ failed:\n" + res.stdout + res.stderr)
    LIB_STAMP.write_text(_lib_hash())
    return out


if __name__ == "__main__":
    import sys
    print(build(force=True, verbose="-v" in sys.argv))
    print(build_host_lib(force=True))
