"""
Token maps of `Kinetics` (reference python/magicsoup/kinetics.py:16-289): every domain token (an index
from the genetics' codon maps) maps to a concrete value -- a Km or Vmax weight, a sign, a Hill
coefficient, or a stoichiometry / transport / effector vector over the signals.  The tables are
plain tensors under the reference's attribute names (`weights`, `signs`, `numbers`, `M`: its tests
overwrite them) with the "undefined" value at token 0; `Kinetics.set_cell_params` hands them to the
device kernel as they are.  Drawn with Python's `random` like the reference's.
"""
from __future__ import annotations

import math
import random

import torch


class LogNormWeightMap:
    """token -> float, log-normal inside weight_range, NaN at token 0 (kinetics.py:40-79)"""

    def __init__(self, max_token: int, weight_range: tuple[float, float], device: str = "cpu", zero_value: float = math.nan):
        lo, hi = min(weight_range), max(weight_range)
        mu, sig = (math.log(lo) + math.log(hi)) / 2, math.log(hi) - math.log(lo)
        vals = [zero_value]
        while len(vals) <= max_token:
            x = math.exp(random.gauss(mu, sig))
            if lo <= x <= hi:
                vals.append(x)
        self.weights = torch.tensor(vals, dtype=torch.float32, device=device)

    def __call__(self, t: torch.Tensor) -> torch.Tensor:
        return self.weights[t]

    def inverse(self) -> dict[float, list[int]]:
        out: dict[float, list[int]] = {}
        for i, v in enumerate(self.weights.cpu().tolist()[1:], start=1):
            out.setdefault(v, []).append(i)
        return out


class SignMap:
    """token -> +1 / -1 with equal chances, 0 at token 0 (kinetics.py:82-103)"""

    def __init__(self, max_token: int, device: str = "cpu", zero_value: int = 0):
        self.signs = torch.tensor([zero_value] + random.choices([1, -1], k=max_token), dtype=torch.int32, device=device)

    def __call__(self, t: torch.Tensor) -> torch.Tensor:
        return self.signs[t]

    def inverse(self) -> dict[bool, list[int]]:
        m = self.signs.cpu()
        return {True: torch.argwhere(m == 1).flatten().tolist(), False: torch.argwhere(m == -1).flatten().tolist()}


class HillMap:
    """token -> 1..5 with chances 52, 26, 13, 6, 3 %, 0 at token 0 (kinetics.py:16-37)"""

    def __init__(self, max_token: int, device: str = "cpu", zero_value: int = 0):
        pool = [5] + 2 * [4] + 4 * [3] + 8 * [2] + 16 * [1]
        self.numbers = torch.tensor([zero_value] + random.choices(pool, k=max_token), dtype=torch.int32, device=device)

    def __call__(self, t: torch.Tensor) -> torch.Tensor:
        return self.numbers[t]

    def inverse(self) -> dict[int, list[int]]:
        m = self.numbers.cpu()
        return {k: torch.argwhere(m == k).flatten().tolist() for k in (1, 3, 5)}


class VectorMap:
    """token -> one of `vectors` (each equally likely), the zero vector at token 0 (kinetics.py:106-150)"""

    def __init__(self, max_token: int, n_signals: int, vectors: list[list[int]], device: str = "cpu", zero_value: int = 0):
        M = torch.full((max_token + 1, n_signals), zero_value, dtype=torch.int32)
        if vectors:
            if any(len(v) != n_signals for v in vectors):
                raise ValueError(f"Not all vectors have length of signal_size={n_signals}")
            if len(vectors) > max_token:
                raise ValueError(f"There are max_token={max_token} and {len(vectors)} vectors. It is not possible to map all vectors")
            if any(all(x == 0 for x in v) for v in vectors):
                raise ValueError("At least one vector includes only zeros. Each vector should contain at least one non-zero value.")
            pick = random.choices(range(len(vectors)), k=max_token)
            M[1:] = torch.tensor(vectors, dtype=torch.int32)[pick]
        self.M = M.to(device)

    def __call__(self, t: torch.Tensor) -> torch.Tensor:
        return self.M[t]


def reaction_vectors(mol_2_idx: dict, reactions: list, n_signals: int) -> list[list[int]]:
    """stoichiometry of every reaction over the intracellular signals (kinetics.py:170-178)"""
    out = []
    for substrates, products in reactions:
        v = [0] * n_signals
        for mol in substrates:
            v[mol_2_idx[mol]] -= 1
        for mol in products:
            v[mol_2_idx[mol]] += 1
        out.append(v)
    return out


def transport_vectors(n_molecules: int) -> list[list[int]]:
    """-1 inside, +1 outside for every molecule (kinetics.py:222-226)"""
    out = []
    for i in range(n_molecules):
        v = [0] * (2 * n_molecules)
        v[i], v[i + n_molecules] = -1, 1
        out.append(v)
    return out


def effector_vectors(n_molecules: int) -> list[list[int]]:
    """one-hot over all 2 * n_molecules signals (kinetics.py:262-266)"""
    n = 2 * n_molecules
    return [[1 if i == j else 0 for j in range(n)] for i in range(n)]
