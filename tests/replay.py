"""
Helpers to replay the recorded reference calls of tests/golden/ref_tests_kinetics.npz
(written by tests/golden/make_golden.py) against any backend.
"""
from __future__ import annotations

from dataclasses import dataclass
from pathlib import Path

import numpy as np

GOLDEN = Path(__file__).resolve().parent / "golden"
PARAMS = ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")


@dataclass
class Record:
    test: str
    method: str
    idx: int
    args: dict
    params: dict
    outs: list
    depth: int
    eq_mocked: bool

    @property
    def id(self) -> str:
        return f"{self.test}-{self.method}-{self.idx}"


def load_records(method: str | None = None) -> list[Record]:
    z = np.load(GOLDEN / "ref_tests_kinetics.npz")
    groups: dict[tuple, dict] = {}
    for key in z.files:
        test, meth, idx, name = key.split("|")
        groups.setdefault((test, meth, int(idx)), {})[name] = z[key]
    recs = []
    for (test, meth, idx), d in sorted(groups.items()):
        if method is not None and meth != method:
            continue
        outs = [d[k] for k in sorted(k for k in d if k.startswith("out_"))]
        recs.append(Record(
            test=test, method=meth, idx=idx,
            args={k[4:]: v for k, v in d.items() if k.startswith("arg_")},
            params={k[4:]: v for k, v in d.items() if k.startswith("par_")},
            outs=outs, depth=int(d["meta"][0]), eq_mocked=bool(d["meta"][1]),
        ))
    return recs


def full_params(rec: Record, c: int, p: int, s: int) -> dict:
    """
    The reference's tests often set only the tensors a helper reads (e.g. Nf/Nb/Ke for
    _get_equilibrium_adjusted_x) and leave the rest at their (0,0[,s]) initial shape.
    Unset ones are filled with zeros of the right shape; they are not read by that call.
    """
    out = {}
    for k in PARAMS:
        t = rec.params[k]
        want = (c, p, s) if k in ("N", "Nf", "Nb", "A", "Kmr") else (c, p)
        if t.shape != want:
            t = np.zeros(want, dtype=t.dtype)
        out[k] = t
    return out


def rel_err(a: np.ndarray, b: np.ndarray, floor: float = 1e-6) -> np.ndarray:
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    with np.errstate(invalid="ignore"):
        d = np.abs(a - b) / np.maximum(np.abs(b), floor)
    d[(a == b) | (np.isnan(a) & np.isnan(b))] = 0.0
    return d


def mismatch_fraction(a, b, tol=1e-4) -> float:
    return float((rel_err(a, b) > tol).mean())


SYNTHETIC_TAGS = ("c1", "wl", "two", "c4", "c4e", "c3", "s82")


def velocity_bound(ref_v: np.ndarray, adj_vmax: np.ndarray) -> np.ndarray:
    """
    SURVEY.md 8(c) P1 for V, stated so that it holds at every tile shape: relative 2e-5 of |V|
    plus 2.5e-7 of the protein's trimmed Vmax (4 float32 ulps of Vmax').  V = a_cat * Vmax' *
    a_reg with a_cat = (kf - kb) / (1 + kf + kb) in (-1, 1): near equilibrium kf - kb cancels,
    so the error of V is absolute in units of Vmax', not relative to |V|.  Measured on the
    fixtures of this directory: max 0.9e-7 * Vmax' (oracle f32 vs reference), 0.65e-7 * Vmax' for
    the reference itself against float64.  (Round 1 used 2e-5 * max(|V|, 1e-3), which 9 of 65,536
    elements of the p=64, s=32 fixture exceed with |V| < 1e-6 beside Vmax' ~ 1.)
    """
    return 2e-5 * np.abs(ref_v) + 2.5e-7 * np.abs(adj_vmax)
