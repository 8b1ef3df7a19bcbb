"""point mutations and recombinations on genome strings (python/magicsoup/mutations.py:4-51)"""
from __future__ import annotations

from magicsoup_b200._hostlib import lib as _lib


def point_mutations(seqs: list[str], p: float = 1e-6, p_indel: float = 0.4, p_del: float = 0.66) -> list[tuple[str, int]]:
    """mutated sequences with their index in seqs; only sequences that changed (mutations.py:4-27)"""
    if len(seqs) == 0:
        return []
    return _lib().point_mutations(seqs, p, p_indel, p_del)


def recombinations(seq_pairs: list[tuple[str, str]], p: float = 1e-7) -> list[tuple[str, str, int]]:
    """recombined pairs with their index in seq_pairs; only pairs that changed (mutations.py:30-51)"""
    if len(seq_pairs) == 0:
        return []
    return _lib().recombinations(seq_pairs, p)
