"""
Helpers with the reference's names and results (python/magicsoup/util.py:1-125): genome
generators and codon enumerations (pure Python, init-time only) and the map helpers that call
the native extension.
"""
from __future__ import annotations

import math
import random
import string
from itertools import product
from typing import Iterable

from magicsoup_b200._hostlib import lib as _lib
from magicsoup_b200.constants import ALL_NTS, CODON_SIZE


def round_down(d: float, to: int = 3) -> int:
    return math.floor(d / to) * to


def closest_value(values: Iterable[float], key: float) -> float:
    return min(values, key=lambda d: abs(d - key))


def randstr(n: int = 12) -> str:
    return "".join(random.choices(string.ascii_uppercase + string.ascii_lowercase + string.digits, k=n))


def random_genome(s: int = 500, excl: list[str] | None = None) -> str:
    """random nucleotide sequence of length s without the sequences in excl (util.py:33-61)"""
    out = "".join(random.choices(ALL_NTS, k=s))
    if excl:
        def strip(seq: str) -> str:
            for bad in excl:
                seq = "".join(seq.split(bad))
            return seq

        out = strip(out)
        while len(out) != s:
            out = strip(out + random_genome(s=s - len(out)))
    return out


def variants(seq: str) -> list[str]:
    """all sequences of a template with N (any), R (A, G), Y (C, T), in the reference's order (util.py:64-87)"""
    pools = {"N": ("T", "C", "G", "A"), "R": ("A", "G"), "Y": ("C", "T")}
    out = [seq]
    for char in ("N", "R", "Y"):
        nxt = []
        for s in out:
            idxs = [i for i, c in enumerate(s) if c == char]
            for combo in product(*([pools[char]] * len(idxs))):
                chars = list(s)
                for i, c in zip(idxs, combo):
                    chars[i] = c
                nxt.append("".join(chars))
        out = nxt
    return out


def codons(n: int, excl_codons: list[str] | None = None) -> list[str]:
    """all sequences of n codons, optionally without those that contain a codon of excl_codons (util.py:90-109)"""
    seqs = variants("N" * n * CODON_SIZE)
    if excl_codons is None:
        return seqs
    excl = set(excl_codons)
    return [s for s in seqs if not any(s[i * CODON_SIZE:(i + 1) * CODON_SIZE] in excl for i in range(n))]


def dist_1d(a: int, b: int, m: int) -> int:
    return _lib().dist_1d(a, b, m)


def free_moores_nghbhd(x: int, y: int, positions: list[tuple[int, int]], map_size: int) -> list[tuple[int, int]]:
    return _lib().free_moores_nghbhd(x, y, positions, map_size)
