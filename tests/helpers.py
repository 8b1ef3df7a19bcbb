"""shared helpers of the GPU parity tests"""
from __future__ import annotations

import numpy as np
import torch

from magicsoup_b200.synthetic import Population, make_population
from oracle import oracle as O

PARAMS = ("N", "Nf", "Nb", "A", "Kmr", "Ke", "Kmf", "Kmb", "Vmax")


def oracle_params(pop: Population) -> O.Params:
    return O.Params(*[t.cpu().numpy() for t in pop.params()])


def install(kin, pop: Population, device="cuda"):
    for k in PARAMS:
        setattr(kin, k, getattr(pop, k).to(device))
    return kin


def bits_equal(a, b) -> bool:
    a = np.ascontiguousarray(a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else a, dtype=np.float32)
    b = np.ascontiguousarray(b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else b, dtype=np.float32)
    if a.shape != b.shape:
        return False
    ai, bi = a.view(np.uint32), b.view(np.uint32)
    same = (ai == bi) | (np.isnan(a) & np.isnan(b))  # bit patterns: +0 and -0 differ
    return bool(same.all())


def describe_mismatch(a, b, name="") -> str:
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    b = b.detach().cpu().numpy() if isinstance(b, torch.Tensor) else np.asarray(b)
    if a.shape != b.shape:
        return f"{name}: shape {a.shape} vs {b.shape}"
    bad = ~((a == b) | (np.isnan(a) & np.isnan(b)))
    if not bad.any():
        return f"{name}: identical"
    idx = np.argwhere(bad)
    with np.errstate(all="ignore"):
        rel = np.abs(a - b) / np.maximum(np.abs(b), 1e-30)
    first = tuple(idx[0])
    rows = np.unique(idx[:, 0])
    return (f"{name}: {bad.sum()} of {bad.size} elements differ in {len(rows)} rows; first at {first}: "
            f"{a[first]!r} vs {b[first]!r}; max rel {np.nanmax(rel[bad]):.3e}; rows {rows[:8].tolist()}")


def dense_adversarial(c, p, s, seed, device="cpu") -> Population:
    """the reference's slow robustness tests' parameter style (tests/slow/test_kinetics.py:130-160)"""
    g = torch.Generator().manual_seed(seed)
    N = torch.randint(-8, 9, (c, p, s), generator=g).int()
    Nf = torch.where(N < 0, -N, 0)
    Nb = torch.where(N > 0, N, 0)
    Vmax = torch.randn(c, p, generator=g).abs().clamp(max=1.0) * 100
    A = torch.randint(-5, 5, (c, p, s), generator=g).int()
    Ke = torch.randn(c, p, generator=g) * 100
    Kmf = torch.randn(c, p, generator=g).abs().clamp(0.001)
    Kmb = (Kmf * Ke).clamp(0.001)
    Kmr = torch.randn(c, p, s, generator=g).abs().clamp(0.001)
    m = s // 2
    side = int(np.ceil(np.sqrt(c))) + 1
    pix = torch.randperm(side * side, generator=g)[:c]
    pop = Population(N=N, Nf=Nf, Nb=Nb, A=A, Kmr=Kmr, Ke=Kmb / Kmf, Kmf=Kmf, Kmb=Kmb, Vmax=Vmax,
                     cell_positions=torch.stack([pix // side, pix % side], 1).int(),
                     cell_molecules=torch.randn(c, m, generator=g).abs().clamp(max=1.0) * 100,
                     molecule_map=torch.randn(m, side, side, generator=g).abs().clamp(max=1.0) * 100,
                     diffusivities=[0.1] * m, permeabilities=[0.1] * m)
    return pop


__all__ = ["make_population", "oracle_params", "install", "bits_equal", "describe_mismatch",
           "dense_adversarial", "PARAMS"]
