"""
numpy/ctypes front end of the CPU oracle (oracle/ms_oracle.c).

TEST INFRASTRUCTURE, NOT PRODUCT CODE: only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs import this module.  The product
(magicsoup_b200/) never does and fails loudly without its CUDA library.

Function names follow the reference (python/magicsoup/kinetics.py:725-918,
python/magicsoup/world.py:610-665, 935-984).  All arrays are numpy, row-major,
shaped like the reference's tensors: X (c,s); N,Nf,Nb,A (c,p,s) int32;
Kmr (c,p,s); Ke,Kmf,Kmb,Vmax (c,p).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIBS: dict[str, C.CDLL] = {}


def build(force: bool = False) -> None:
    """compile libms_oracle_f32.so / _f64.so next to this file (gcc + make)"""
    if force:
        subprocess.run(["make", "-C", str(_HERE), "clean"], check=True, capture_output=True)
    subprocess.run(["make", "-C", str(_HERE), "all"], check=True, capture_output=True)


def _lib(dtype) -> C.CDLL:
    key = "f32" if np.dtype(dtype) == np.float32 else "f64"
    if key not in _LIBS:
        path = _HERE / f"libms_oracle_{key}.so"
        src = _HERE / "ms_oracle.c"
        if not path.exists() or path.stat().st_mtime < src.stat().st_mtime:
            build()
        _LIBS[key] = C.CDLL(str(path))
    return _LIBS[key]


def _p(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _r(a, dtype) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a), dtype=dtype)


def _i(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a), dtype=np.int32)


def _i64(x) -> C.c_int64:
    return C.c_int64(int(x))


class Params:
    """the nine Kinetics parameter tensors (kinetics.py:399-409) as numpy arrays"""

    def __init__(self, N, Nf, Nb, A, Kmr, Ke, Kmf, Kmb, Vmax, dtype=np.float32):
        self.dtype = np.dtype(dtype)
        self.N, self.Nf, self.Nb, self.A = _i(N), _i(Nf), _i(Nb), _i(A)
        self.Kmr = _r(Kmr, dtype)
        self.Ke, self.Kmf, self.Kmb, self.Vmax = (_r(t, dtype) for t in (Ke, Kmf, Kmb, Vmax))
        self.c, self.p, self.s = self.N.shape

    def astype(self, dtype) -> "Params":
        return Params(self.N, self.Nf, self.Nb, self.A, self.Kmr, self.Ke, self.Kmf,
                      self.Kmb, self.Vmax, dtype=dtype)

    def _args(self):
        return [_p(t) for t in (self.N, self.Nf, self.Nb, self.A, self.Kmr, self.Ke,
                                self.Kmf, self.Kmb)]


def multiply_signals(X, N, dtype=np.float32):
    """kinetics.py:894-918 -> (xx (c,p), involved (c,p) bool)"""
    X, N = _r(X, dtype), _i(N)
    c, p, s = N.shape
    xx = np.empty((c, p), dtype=dtype)
    inv = np.empty((c, p), dtype=np.uint8)
    _lib(dtype).ms_oracle_multiply_signals(_p(X), _p(N), _i64(c), _i64(p), _i64(s), _p(xx), _p(inv))
    return xx, inv.astype(bool)


def get_quotient(X, Nf, Nb, dtype=np.float32):
    """kinetics.py:881-892 -> Q (c,p)"""
    X, Nf, Nb = _r(X, dtype), _i(Nf), _i(Nb)
    c, p, s = Nf.shape
    Q = np.empty((c, p), dtype=dtype)
    _lib(dtype).ms_oracle_quotient(_p(X), _p(Nf), _p(Nb), _i64(c), _i64(p), _i64(s), _p(Q))
    return Q


def get_velocities(P: Params, X, Vmax):
    """kinetics.py:771-806 -> V (c,p); Vmax is the already trimmed adj_vmax"""
    X, Vmax = _r(X, P.dtype), _r(Vmax, P.dtype)
    V = np.empty((P.c, P.p), dtype=P.dtype)
    _lib(P.dtype).ms_oracle_velocities(_p(X), _p(Vmax), _p(P.Nf), _p(P.Nb), _p(P.A), _p(P.Kmr),
                                       _p(P.Kmf), _p(P.Kmb), _i64(P.c), _i64(P.p), _i64(P.s), _p(V))
    return V


def get_negative_adjusted_nv(NV, X, dtype=np.float32):
    """kinetics.py:861-879 -> NV_adj (c,p,s)"""
    NV, X = _r(NV, dtype), _r(X, dtype)
    c, p, s = NV.shape
    out = np.empty_like(NV)
    _lib(dtype).ms_oracle_negative_adjusted_nv(_p(NV), _p(X), _i64(c), _i64(p), _i64(s), _p(out))
    return out


def get_equilibrium_adjusted_x(X0, X1, NV, V, Nf, Nb, Ke, dtype=np.float32, n_threads=0):
    """kinetics.py:808-859 -> (X1_adj (c,s), number of F updates applied)"""
    X0, X1, NV, V, Ke = (_r(t, dtype) for t in (X0, X1, NV, V, Ke))
    Nf, Nb = _i(Nf), _i(Nb)
    c, p, s = NV.shape
    out = np.empty_like(X1)
    done = C.c_int(0)
    rc = _lib(dtype).ms_oracle_equilibrium_adjusted_x(
        _p(X0), _p(X1), _p(NV), _p(V), _p(Nf), _p(Nb), _p(Ke), _i64(c), _i64(p), _i64(s),
        _p(out), C.byref(done), C.c_int(n_threads))
    assert rc == 0
    return out, done.value


def integrate_signals_part(P: Params, adj_vmax, X0, max_eq_iters=4, n_threads=0):
    """kinetics.py:753-769 -> (X1 (c,s), V (c,p), number of F updates applied)"""
    X0, adj = _r(X0, P.dtype), _r(adj_vmax, P.dtype)
    X1 = np.empty_like(X0)
    V = np.empty((P.c, P.p), dtype=P.dtype)
    done = C.c_int(0)
    rc = _lib(P.dtype).ms_oracle_integrate_part(
        *P._args(), _p(adj), _i64(P.c), _i64(P.p), _i64(P.s), _p(X0), _p(X1),
        C.c_int(max_eq_iters), C.byref(done), _p(V), C.c_int(n_threads))
    assert rc == 0
    return X1, V, done.value


def integrate_signals(P: Params, X, n_threads=0, return_eq_done=False):
    """kinetics.py:725-751 -> X (c,s) after the three trimmed parts"""
    X = _r(X, P.dtype)
    out = np.empty_like(X)
    done = (C.c_int * 3)()
    rc = _lib(P.dtype).ms_oracle_integrate_signals(
        *P._args(), _p(P.Vmax), _i64(P.c), _i64(P.p), _i64(P.s), _p(X), _p(out), done,
        C.c_int(n_threads))
    assert rc == 0
    return (out, list(done)) if return_eq_done else out


def integrate_signals_iters(P: Params, X, eq_iters, n_threads=0):
    """
    CPU mirror of the contract of the CUDA entry point ms_integrate_signals: applies
    eq_iters[q] adjustment iterations per part and returns (X_out, flags[24]).
    """
    X = _r(X, P.dtype)
    out = np.empty_like(X)
    flags = np.zeros(24, dtype=np.uint32)
    eq = (C.c_int * 3)(*[int(v) for v in eq_iters])
    rc = _lib(P.dtype).ms_oracle_integrate_signals_iters(
        *P._args(), _p(P.Vmax), _i64(P.c), _i64(P.p), _i64(P.s), _p(X), _p(out), eq, _p(flags),
        C.c_int(n_threads))
    assert rc == 0
    return out, flags


def diffuse_stencil(map_slab, halo_top, halo_bottom, a, b, dtype=np.float32):
    """slab stencil -> (out (m,H,W), sums int64[4*m]); a, b float32 weight vectors"""
    mm = _r(map_slab, dtype)
    m, H, W = mm.shape
    top = None if halo_top is None else _r(halo_top, dtype)
    bot = None if halo_bottom is None else _r(halo_bottom, dtype)
    out = np.empty_like(mm)
    sums = np.zeros(4 * m, dtype=np.int64)
    rc = _lib(dtype).ms_oracle_diffuse_stencil(_p(mm), _p(top), _p(bot), _p(_r(a, dtype)), _p(_r(b, dtype)),
                                               _i64(m), _i64(H), _i64(W), _p(out), _p(sums))
    assert rc == 0
    return out, sums


def diffuse_correct(out, sums, n_pxls_total, dtype=np.float32):
    """second pass of the slab diffusion -> new map slab"""
    o = _r(out, dtype)
    m, H, W = o.shape
    res = np.empty_like(o)
    sums = np.ascontiguousarray(sums, dtype=np.int64)
    rc = _lib(dtype).ms_oracle_diffuse_correct(_p(o), _p(res), _p(sums), _i64(m), _i64(H), _i64(W),
                                               _i64(n_pxls_total))
    assert rc == 0
    return res


def enzymatic_activity(P: Params, positions, cell_molecules, molecule_map, n_threads=0):
    """world.py:610-625; returns new (cell_molecules, molecule_map), inputs untouched"""
    pos = _i(positions)
    cm = _r(cell_molecules, P.dtype).copy()
    mm = _r(molecule_map, P.dtype).copy()
    m, H, W = mm.shape
    assert P.s == 2 * m
    rc = _lib(P.dtype).ms_oracle_enzymatic_activity(
        *P._args(), _p(P.Vmax), _i64(P.c), _i64(P.p), _i64(m), _p(pos), _p(cm), _p(mm),
        _i64(H), _i64(W), None, C.c_int(n_threads))
    assert rc == 0
    return cm, mm


def diffusion_weights(diffusivity: float) -> tuple[float, float]:
    """world.py:948-964 _get_diffuse: 3x3 kernel weights (a = neighbours, b = centre)"""
    r = abs(float(diffusivity))
    r = min(r, 1.0)
    if r == 0.0:
        return 0.0, 1.0
    d = 1 / r
    a = 1 / (d + 8)
    b = d * a
    b = b + 1.0 - (8 * a + b)
    return a, b


def permeation_factor(permeability: float) -> float:
    """world.py:935-946 _get_permeate"""
    r = min(abs(float(permeability)), 1.0)
    if r == 0.0:
        return 0.0
    return 1 / (1 / r + 1)


def diffuse(molecule_map, diffusivities, dtype=np.float32):
    """world.py:638-649 for all molecules; returns the new map"""
    mm = _r(molecule_map, dtype).copy()
    m, H, W = mm.shape
    ab = [diffusion_weights(d) for d in diffusivities]
    # the reference stores the weights as a float32 conv kernel (world.py:967-972)
    a = _r(np.array([x[0] for x in ab], dtype=np.float32), dtype)
    b = _r(np.array([x[1] for x in ab], dtype=np.float32), dtype)
    rc = _lib(dtype).ms_oracle_diffuse(_p(mm), _p(a), _p(b), _i64(m), _i64(H), _i64(W), C.c_int(0))
    assert rc == 0
    return mm


def permeate(cell_molecules, molecule_map, positions, permeabilities, dtype=np.float32):
    """world.py:651-665; returns new (cell_molecules, molecule_map)"""
    cm = _r(cell_molecules, dtype).copy()
    mm = _r(molecule_map, dtype).copy()
    pos = _i(positions)
    m, H, W = mm.shape
    # python float * float32 tensor: the scalar is rounded to float32 first
    q = _r(np.array([permeation_factor(d) for d in permeabilities], dtype=np.float32), dtype)
    rc = _lib(dtype).ms_oracle_permeate(_p(cm), _p(mm), _p(pos), _p(q), _i64(cm.shape[0]),
                                        _i64(m), _i64(H), _i64(W))
    assert rc == 0
    return cm, mm


def diffuse_molecules(cell_molecules, molecule_map, positions, diffusivities, permeabilities,
                      dtype=np.float32):
    """world.py:627-665: diffusion of every plane, then membrane permeation"""
    mm = diffuse(molecule_map, diffusivities, dtype=dtype)
    if len(cell_molecules) == 0:
        return _r(cell_molecules, dtype).copy(), mm
    return permeate(cell_molecules, mm, positions, permeabilities, dtype=dtype)
